"""Callers either side of the compare step: test-time mask policies and the
test-phase refine iteration, against the numpy lifts of the reference."""
import numpy as np
import pytest

from oracle import updater_np, zoom_np

pytestmark = pytest.mark.gpu


def _masks(rng, B, H, W):
    m = np.zeros((B, 1, H, W), np.float32)
    for b in range(B):
        y0, x0 = rng.integers(0, H - 60), rng.integers(0, W - 60)
        h, w = rng.integers(2, 60), rng.integers(2, 60)
        blob = (rng.random((h, w)) > 0.4).astype(np.float32)
        blob[0, 0] = 1  # never empty
        m[b, 0, y0:y0 + h, x0:x0 + w] = blob
    return m


def test_mask_box_matches_update_data_batch(cuda_device):
    import torch

    from mx_deepim_b200 import pipeline

    rng = np.random.default_rng(3)
    B, H, W = 5, 480, 640
    m = _masks(rng, B, H, W)
    m[1, 0] = 0
    m[1, 0, 0, 0] = 1          # single pixel in the corner: empty slice
    m[2, 0] = 0
    m[2, 0, H - 1, W - 4:] = 1  # touches the last row / column
    m[3, 0] = 1                 # full frame
    d = torch.from_numpy(m).to(cuda_device)
    ws = torch.zeros((B, 5), dtype=torch.int32, device=cuda_device)
    for rep in range(2):  # second call checks that the workspace cleaned itself
        out, bbox = pipeline.mask_box(d, workspace=ws)
        torch.cuda.synchronize()
        out, bbox = out.cpu().numpy(), bbox.cpu().numpy()
        for b in range(B):
            for policy, obs, rend in (("box_rendered", np.zeros((H, W)), m[b, 0]),
                                      ("box_observed", m[b, 0], np.zeros((H, W)))):
                ref = updater_np.update_mask_observed(policy, obs, rend)
                assert np.array_equal(out[b, 0], ref.astype(np.float32)), (b, policy)
            ys, xs = np.nonzero(m[b, 0])
            assert list(bbox[b]) == [xs.min(), xs.max(), ys.min(), ys.max()]
        assert int(ws.abs().sum()) == 0


def test_mask_box_empty_mask(cuda_device):
    import torch

    from mx_deepim_b200 import pipeline

    d = torch.zeros((2, 1, 480, 640), device=cuda_device)
    d[1, 0, 100:120, 200:230] = 1
    out, bbox = pipeline.mask_box(d)
    torch.cuda.synchronize()
    bbox = bbox.cpu().numpy()
    assert bbox[0, 0] > bbox[0, 1] and bbox[0, 2] > bbox[0, 3]  # the reference raises here
    assert float(out[0].abs().sum()) == 0
    with pytest.raises(Exception):
        updater_np.update_mask_observed("box_observed", np.zeros((480, 640)), None)
    assert float(out[1].sum()) == 19 * 29


def test_test_phase_iteration(cuda_device):
    """tester.py:520-594 + data_pair.update_data_batch + the test graph's zoom ops."""
    import torch

    from mx_deepim_b200 import pipeline, synth

    rng = np.random.default_rng(21)
    meshes = [synth.make_mesh(rng, level=3, kind=0)]
    B = 2
    loop = pipeline.RefineLoop(meshes, B, cuda_device, seed=1)
    host = pipeline.make_refine_batch(meshes, B, 77, renderer=loop.renderer)
    loop.set_batch(host)
    loop.iterate(0)  # leaves a rendered mask of the current pose in the batch
    torch.cuda.synchronize()
    prev_mask = loop.mask_rendered.cpu().numpy().copy()
    prev_pose = loop.pose[loop.cur].cpu().numpy().copy()
    rot = loop.rot_est.clone()
    trans = loop.trans_est.clone()
    for policy in ("box_rendered", "init"):
        loop.iterate_test(rot, trans, update_mask=policy)
        torch.cuda.synchronize()
        pose = loop.pose[loop.cur].cpu().numpy()
        ref_pose, _, _, _ = updater_np.pose_update(prev_pose, rot.cpu().numpy(), trans.cpu().numpy(),
                                                   prev_pose, synth.K_LINEMOD, np.zeros(3, np.float32),
                                                   np.ones(3, np.float32))
        np.testing.assert_allclose(pose, ref_pose, rtol=2e-6, atol=2e-7)
        depth = loop.depth_rendered.cpu().numpy()
        if policy == "box_rendered":
            mo = np.stack([updater_np.update_mask_observed("box_rendered", None, prev_mask[b, 0])
                           for b in range(B)])[:, None].astype(np.float32)
        else:
            mo = host["mask_observed"]
        # test graph: mask_gt_observed = mask_observed (deepIM_flownet.py:781)
        zm = zoom_np.zoom_mask_forward(mo, mo, depth, pose, synth.K_LINEMOD)
        assert np.array_equal(loop.step.zoom_factor.cpu().numpy(), zm[3])
        assert (loop.step.zoom_mask_observed.cpu().numpy() != zm[0]).mean() < 1e-5
        assert (loop.step.zoom_mask_rendered.cpu().numpy() != zm[2]).mean() < 1e-5
        zi = zoom_np.zoom_image_with_factor_forward(zm[3], host["image_observed"],
                                                    loop.image_rendered.cpu().numpy(),
                                                    synth.PIXEL_MEANS[::-1])
        assert np.abs(loop.step.zoom_image_observed.cpu().numpy() - zi[0]).max() <= 1e-4
        assert np.abs(loop.step.zoom_image_rendered.cpu().numpy() - zi[1]).max() <= 1e-4
        prev_mask = loop.mask_rendered.cpu().numpy().copy()
        prev_pose = pose.copy()
