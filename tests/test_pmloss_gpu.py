"""Transform3D + point-matching loss on the GPU vs oracle/transform3d_np.py."""
import numpy as np
import pytest

import helpers
from oracle import transform3d_np as t3d

pytestmark = pytest.mark.gpu
MEANS = np.zeros(3, np.float32)
STDS = np.ones(3, np.float32)


def _t(a, dev):
    import torch

    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


@pytest.mark.parametrize("rot_coord", ["CAMERA", "MODEL", "CAMERA_NEW", "NAIVE"])
def test_transform3d_forward_backward(cuda_device, rot_coord):
    import torch

    from mx_deepim_b200 import operator, operator_py  # noqa: F401

    d = helpers.make_points_inputs(8, 21, num_points=3000, num_keep=2500)
    ref = t3d.forward(d["points"], d["rotation"], d["translation"], d["pose_src"], MEANS, STDS,
                      rot_coord)
    h = operator.Custom(op_type="Transform3D", point_cloud=_t(d["points"], cuda_device),
                        rotation=_t(d["rotation"], cuda_device),
                        translation=_t(d["translation"], cuda_device),
                        pose_src=_t(d["pose_src"], cuda_device), T_means=MEANS, T_stds=STDS,
                        rot_coord=rot_coord)
    torch.cuda.synchronize()
    got = h.outputs[0].cpu().numpy()
    assert np.abs(got - ref).max() < 1e-5  # reference's own bar is 1e-4 (transform3d.py:487-493)
    rng = np.random.default_rng(5)
    og = rng.normal(size=ref.shape).astype(np.float32)
    g_rot, g_tr = t3d.backward(og, d["points"], d["rotation"], d["translation"], d["pose_src"],
                               MEANS, STDS, rot_coord)
    grads = h.backward([_t(og, cuda_device)])
    torch.cuda.synchronize()
    assert float(grads[0].abs().sum()) == 0.0 and float(grads[3].abs().sum()) == 0.0
    np.testing.assert_allclose(grads[1].cpu().numpy(), g_rot, rtol=2e-4, atol=2e-4)
    np.testing.assert_allclose(grads[2].cpu().numpy(), g_tr, rtol=2e-4, atol=2e-4)


@pytest.mark.parametrize("loss_type", ["L1", "L2", "smooth_L1"])
def test_pm_loss_fused(cuda_device, loss_type):
    import torch

    from mx_deepim_b200.operator_py.transform3d import (point_matching_loss,
                                                        point_matching_loss_backward)

    d = helpers.make_points_inputs(16, 22, num_points=3000, num_keep=2800)
    args = [d[k] for k in ("points", "points_obs", "weights", "rotation", "translation", "pose_src")]
    ref_map, ref_est, ref_sum = t3d.pm_loss_forward(*args, MEANS, STDS, "CAMERA", loss_type, 0.1, 1.0)
    targs = [_t(a, cuda_device) for a in args]
    loss_map, loss_sum, est = point_matching_loss(*targs, T_means=MEANS, T_stds=STDS,
                                                  rot_coord="CAMERA", loss_type=loss_type,
                                                  norm_term=0.1, sl1_scalar=1.0, want_est=True)
    torch.cuda.synchronize()
    assert np.abs(est.cpu().numpy() - ref_est).max() < 1e-5
    np.testing.assert_allclose(loss_map.cpu().numpy(), ref_map, rtol=1e-4, atol=1e-4)
    np.testing.assert_allclose(loss_sum.cpu().numpy(), ref_sum, rtol=1e-4)
    # padded points carry zero weight
    assert not loss_map.cpu().numpy()[:, :, 2800:].any()
    gs = 0.1 / 3000  # LW_PM / NUM_3D_SAMPLE (deepIM_flownet.py:493-497)
    g_rot, g_tr = t3d.pm_loss_backward(*args, MEANS, STDS, gs, "CAMERA", loss_type, 0.1, 1.0)
    gr, gt = point_matching_loss_backward(*targs, gs, T_means=MEANS, T_stds=STDS,
                                          rot_coord="CAMERA", loss_type=loss_type, norm_term=0.1,
                                          sl1_scalar=1.0)
    torch.cuda.synchronize()
    np.testing.assert_allclose(gr.cpu().numpy(), g_rot, rtol=5e-4, atol=1e-6)
    np.testing.assert_allclose(gt.cpu().numpy(), g_tr, rtol=5e-4, atol=1e-6)


def test_pm_loss_degenerate_quaternion(cuda_device):
    """Non-unit quaternion: identity rotation forward (tol 1e-2), zero rotation grad (tol 1e-4)."""
    import torch

    from mx_deepim_b200.operator_py.transform3d import (point_matching_loss,
                                                        point_matching_loss_backward)

    d = helpers.make_points_inputs(2, 23, num_points=256)
    d["rotation"][0] *= 1.2     # outside 1e-2: identity forward
    d["rotation"][1] *= 1.001   # inside 1e-2 but outside 1e-4: zero rot grad only
    args = [d[k] for k in ("points", "points_obs", "weights", "rotation", "translation", "pose_src")]
    ref_map, ref_est, _ = t3d.pm_loss_forward(*args, MEANS, STDS)
    targs = [_t(a, cuda_device) for a in args]
    _, _, est = point_matching_loss(*targs, T_means=MEANS, T_stds=STDS, want_est=True)
    assert np.abs(est.cpu().numpy() - ref_est).max() < 1e-5
    gr, gt = point_matching_loss_backward(*targs, 1.0, T_means=MEANS, T_stds=STDS)
    assert float(gr.abs().sum()) == 0.0
    g_rot, g_tr = t3d.pm_loss_backward(*args, MEANS, STDS, 1.0)
    np.testing.assert_allclose(gt.cpu().numpy(), g_tr, rtol=5e-4, atol=1e-5)
