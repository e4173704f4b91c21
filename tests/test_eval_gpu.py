"""Evaluation reductions (pose errors, visibility masks, flow EPE) on the device against the
outputs of the reference's own pose_error.py / visibility.py (tests/golden/eval_reference.npz)
and the numpy lifts in oracle/eval_np.py."""
import os

import numpy as np
import pytest

import helpers
from oracle import eval_np as E

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_pose_errors_match_reference_golden(cuda_device):
    import torch

    from mx_deepim_b200 import evaluation as ev

    g = np.load(os.path.join(GOLD, "eval_reference.npz"))
    inp = helpers.make_eval_inputs(7)
    t = lambda a: torch.from_numpy(a).to(cuda_device)
    out = ev.pose_errors(t(inp["pose_est"]), t(inp["pose_gt"]), t(inp["points"]), helpers.K_LM,
                         n_valid=t(inp["n_valid"]))
    torch.cuda.synchronize()
    out = out.cpu().numpy()
    ref = g["pose_errors"]
    np.testing.assert_allclose(out[:, 0], ref[:, 0], rtol=1e-9)   # add
    np.testing.assert_allclose(out[:, 1], ref[:, 1], rtol=1e-9)   # adi (exact NN == k-d tree NN)
    np.testing.assert_allclose(out[:, 2], ref[:, 2], rtol=1e-6)   # arp_2d (K is float32 on device)
    np.testing.assert_allclose(out[:, 3], ref[:, 3], atol=1e-4)   # re, degrees (logm of a float32 R)
    np.testing.assert_allclose(out[:, 4], ref[:, 4], rtol=1e-9)   # te
    # shared point set + the reference-named single-pair wrappers
    n = int(inp["n_valid"][0])
    pts = inp["points"][0, :, :n]
    out2 = ev.pose_errors(t(inp["pose_est"]), t(inp["pose_gt"]), t(np.ascontiguousarray(pts)),
                          helpers.K_LM, want_adi=False)
    torch.cuda.synchronize()
    assert abs(float(out2[0, 0]) - ref[0, 0]) < 1e-9 and float(out2[0, 1]) == 0.0
    Re, te_, Rg, tg = (inp["pose_est"][0, :, :3], inp["pose_est"][0, :, 3], inp["pose_gt"][0, :, :3],
                       inp["pose_gt"][0, :, 3])
    assert abs(ev.add(Re, te_, Rg, tg, pts.T) - ref[0, 0]) < 1e-9
    assert abs(ev.adi(Re, te_, Rg, tg, pts.T) - ref[0, 1]) < 1e-9
    assert abs(ev.arp_2d(Re, te_, Rg, tg, pts.T, helpers.K_LM) - ref[0, 2]) < 1e-4
    assert abs(ev.re(Re, Rg) - ref[0, 3]) < 1e-4 and abs(ev.te(te_, tg) - ref[0, 4]) < 1e-9
    assert ev.re(np.eye(3), np.eye(3)) == 0.0


def test_visibility_masks_match_reference_golden(cuda_device):
    from mx_deepim_b200 import evaluation as ev

    g = np.load(os.path.join(GOLD, "eval_reference.npz"))
    inp = helpers.make_eval_inputs(7)
    vg = ev.estimate_visib_mask_gt(inp["d_test"], inp["d_gt"], 0.015)
    ve = ev.estimate_visib_mask_est(inp["d_test"], inp["d_est"], vg, 0.015)
    assert np.array_equal(np.packbits(vg), g["visib_gt"])     # bit-exact
    assert np.array_equal(np.packbits(ve), g["visib_est"])


def test_flow_epe_matches_tester(cuda_device):
    import torch

    from mx_deepim_b200 import evaluation as ev

    rng = np.random.default_rng(2)
    H, W = 480, 640
    gt = rng.normal(0, 5, size=(H, W, 2)).astype(np.float32)
    pred = gt + rng.normal(0, 1, size=(H, W, 2)).astype(np.float32)
    depth = np.zeros((H, W), np.float32)
    depth[100:300, 200:500] = 0.8
    visible = np.zeros((H, W))
    visible[120:280, 220:400] = 1
    bg = np.logical_and(visible == 0, depth == 0)  # tester.py:712
    ref = E.calc_EPE_one_pair(pred.astype(np.float64), gt.astype(np.float64), visible, bg)
    got = ev.calc_EPE_one_pair([pred], [[gt, visible, bg]], 0)
    for k in ref:
        np.testing.assert_allclose(got[k], ref[k], rtol=1e-9)
    # batched form deriving bg from the rendered depth
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(cuda_device)
    out = ev.flow_epe(t(pred.transpose(2, 0, 1)[None]), t(gt.transpose(2, 0, 1)[None]),
                      t(visible.astype(np.float32)[None, None]), depth_rendered=t(depth[None, None]))
    torch.cuda.synchronize()
    o = out[0].cpu().numpy()
    np.testing.assert_allclose(o, [ref["epe_all"], ref["num_all"], ref["epe_viz"], ref["num_viz"],
                                   ref["epe_vizbg"], ref["num_vizbg"]], rtol=1e-9)
