"""The on-device refine loop at the BASELINE.json shapes, checked against the oracles on a seeded
SAMPLE of pairs in every iteration:

    cfg3  B = 64,   one ~5k-triangle mesh
    cfg4  B = 256,  8 meshes, occluders in the observed depth (Occlusion-LINEMOD-shaped)
    cfg5  B = 1024, 16 varied 20k-triangle meshes, PM-loss backward + invZoomTrans backward

For >= 8 sampled pairs per iteration the device's refined pose, regression labels, rendered depth /
mask, GT flow + validity, crop box / zoom factor, zoomed planes, point-matching loss (on the
PRE-update pose, as the reference's network pass computes it) and gradients are compared with
oracle/updater_np (numpy restatement of batch_updater_py_multi.py:100-387 + C rasteriser + C
flow_kernel), oracle/zoom_np and oracle/transform3d_np at the north_star bars.  The loop runs with
the sparse frame step and the incremental background (RefineLoop defaults); the final state of
the eager loop is then compared bit for bit with RefineLoop.replay() of the captured loop.
Reference: deepim/core/module.py:1205-1213 -> lib/pair_matching/batch_updater_py_multi.py:100-387."""
import numpy as np
import pytest

from oracle import transform3d_np, updater_np, zoom_np

pytestmark = pytest.mark.gpu

SHAPES = {
    3: dict(B=64, level=4, meshes=1, occlusion=False, backward=False),
    4: dict(B=256, level=4, meshes=8, occlusion=True, backward=False),
    5: dict(B=1024, level=5, meshes=16, occlusion=False, backward=True),
}
N_SAMPLE = 8
ZERO, ONE = np.zeros(3, np.float32), np.ones(3, np.float32)


def _take(t, idx):
    return t[idx].cpu().numpy()


@pytest.mark.parametrize("cfg", [3, 4, 5])
def test_refine_loop_at_baseline_shape(cuda_device, cfg):
    import torch

    from mx_deepim_b200 import pipeline, synth

    sh = SHAPES[cfg]
    B = sh["B"]
    rng = np.random.default_rng(20181804 + cfg)  # bench.py's generator for this config
    meshes = [synth.make_mesh(rng, level=sh["level"]) for _ in range(sh["meshes"])]
    loop = pipeline.RefineLoop(meshes, B, cuda_device, seed=0, backward=sh["backward"])
    host = pipeline.make_refine_batch(meshes, B, 20181804 + cfg, renderer=loop.renderer,
                                      occlusion=sh["occlusion"])
    loop.set_batch(host)
    pm = synth.PIXEL_MEANS[[2, 1, 0]]
    srng = np.random.default_rng(1000 + cfg)
    iters = 4
    for it in range(iters):
        idx_np = np.sort(srng.choice(B, size=N_SAMPLE, replace=False))
        idx = torch.from_numpy(idx_np).to(cuda_device)
        prev = _take(loop.pose[loop.cur], idx)
        loop.iterate(it)
        torch.cuda.synchronize()
        rot_est, trans_est = _take(loop.rot_est, idx), _take(loop.trans_est, idx)
        cls = host["class_index"][idx_np]
        ref = updater_np.updater_forward(meshes, cls, prev, host["tgt_pose"][idx_np], rot_est, trans_est,
                                         host["depth_gt_observed"][idx_np], synth.K_LINEMOD, pm, ZERO, ONE)
        pose = _take(loop.pose[loop.cur], idx)
        np.testing.assert_allclose(pose, ref["src_pose"], rtol=2e-6, atol=2e-7)
        np.testing.assert_allclose(_take(loop.rot_gt, idx), ref["rot"], atol=3e-6)
        np.testing.assert_allclose(_take(loop.trans_gt, idx), ref["trans"], rtol=1e-5, atol=2e-6)
        np.testing.assert_allclose(_take(loop.KT, idx), ref["KT"], rtol=2e-5, atol=2e-4)
        # ---- render + post (a2/a4): mask >= 99.9 %, depth 1e-5 relative
        depth = _take(loop.depth_rendered, idx)
        fg, rfg = depth > 0, ref["depth_rendered"] > 0
        assert fg.sum() > 50 * N_SAMPLE
        assert (fg == rfg).mean() >= 0.999, (cfg, it, (fg == rfg).mean())
        both = fg & rfg
        rel = np.abs(depth[both] - ref["depth_rendered"][both]) / ref["depth_rendered"][both]
        assert (rel <= 1e-5).mean() >= 0.999, (cfg, it, rel.max())
        assert np.array_equal(_take(loop.mask_rendered, idx), (depth > 0.2).astype(np.float32))
        img = _take(loop.image_rendered, idx)
        b3 = np.broadcast_to(both, img.shape)
        assert (np.abs(img[b3] - ref["image_rendered"][b3]) < 1e-3).mean() >= 0.995
        # ---- GT flow (a8): validity >= 99.9 % (bit-exact wherever the depth inputs agree)
        valid = _take(loop.step.valid, idx)
        assert (valid == ref["flow_weights"][:, :1]).mean() >= 0.999
        v2 = (valid > 0) & (ref["flow_weights"][:, :1] > 0)
        fl = _take(loop.step.flow, idx)
        for c in range(2):
            if v2.any():
                assert np.abs(fl[:, c:c + 1][v2] - ref["flow"][:, c:c + 1][v2]).max() < 2e-3
        # ---- zoom (a10/a12/a14/a16) on the device's own intermediates: boxes and factors bit-exact
        zm = zoom_np.zoom_mask_forward(host["mask_observed"][idx_np], host["mask_gt_observed"][idx_np],
                                       depth, pose, synth.K_LINEMOD)
        assert np.array_equal(_take(loop.step.zoom_factor, idx), zm[3]), (cfg, it)
        assert np.array_equal(_take(loop.step.bbox, idx).astype(np.int64), zm[4])
        for k, i in (("zoom_mask_observed", 0), ("zoom_mask_gt_observed", 1), ("zoom_mask_rendered", 2)):
            assert (_take(getattr(loop.step, k), idx) != zm[i]).mean() < 1e-5, k
        zi = zoom_np.zoom_image_with_factor_forward(zm[3], host["image_observed"][idx_np], img,
                                                    synth.PIXEL_MEANS[::-1])
        assert np.abs(_take(loop.step.zoom_image_observed, idx) - zi[0]).max() <= 1e-4
        assert np.abs(_take(loop.step.zoom_image_rendered, idx) - zi[1]).max() <= 1e-4
        zf, zw = zoom_np.zoom_flow_forward(zm[3], fl, np.tile(valid, [1, 2, 1, 1]))
        wx = zm[3][:, 0].reshape(-1, 1, 1, 1)
        assert (np.abs(_take(loop.step.zoom_flow, idx) - zf) * wx).max() <= 1e-4  # 1e-4 px before /wx
        assert (_take(loop.step.zoom_flow_weights, idx) != zw).mean() < 1e-5
        zt = zoom_np.zoom_trans_forward(zm[3], _take(loop.trans_gt, idx))
        assert np.array_equal(_take(loop.step.zoom_trans_gt, idx), zt)
        # ---- point-matching loss of the pass that produced (rot_est, trans_est): PRE-update pose
        args = (host["point_cloud_model"][idx_np], host["point_cloud_observed"][idx_np],
                host["point_cloud_weights"][idx_np], rot_est, trans_est, prev, ZERO, ONE)
        _, _, sums = transform3d_np.pm_loss_forward(*args)
        np.testing.assert_allclose(_take(loop.step.loss_sum, idx), sums, rtol=1e-4)
        if sh["backward"]:
            g_rot, g_tr = transform3d_np.pm_loss_backward(*args, 1.0 / synth.NUM_3D_SAMPLE)
            # L1 loss: d|x|/dx = sign(x) FLIPS for a coordinate whose residual is within float32
            # rounding of zero, and every flip moves a gradient component by up to
            # 2 * grad_scale / norm * |d est / d parameter| -- as the poses converge (iteration 2, 3)
            # more and more residuals sit there.  Count them in the oracle and allow that much.
            est = transform3d_np.forward(args[0], rot_est, trans_est, prev, ZERO, ONE)
            x = (est - args[1]) / np.float32(0.1)
            near = ((np.abs(x) < 2e-5) & (args[2] > 0)).reshape(N_SAMPLE, -1).sum(axis=1, keepdims=True)
            flip = 2.0 * (1.0 / synth.NUM_3D_SAMPLE) / 0.1 * 4.0
            for got, want in ((_take(loop.step.grad_rot, idx), g_rot), (_take(loop.step.grad_trans, idx), g_tr)):
                slack = near * flip + 1e-3 * np.abs(want).max(axis=1, keepdims=True) + 1e-6
                assert (np.abs(got - want) <= 5e-4 * np.abs(want) + slack).all(), (cfg, it, near.ravel())
    loop.check_status()
    # the poses converged towards the targets under the damped-GT stand-in
    err0 = np.abs(host["src_pose"] - host["tgt_pose"]).max(axis=(1, 2))
    err4 = np.abs(loop.pose[loop.cur].cpu().numpy() - host["tgt_pose"]).max(axis=(1, 2))
    assert np.median(err4) < 0.35 * np.median(err0)
    def direct_grads():
        """rc_pm_loss_bwd_dev called directly on the loop's final (rot_est, trans_est, pre-update
        pose): what the fused step's backward leg must hold, whichever stream / graph ran it."""
        from mx_deepim_b200 import _lib

        stp, o = loop.step, loop.obs
        gr, gt = torch.empty_like(stp.grad_rot), torch.empty_like(stp.grad_trans)
        rc = _lib.load().rc_pm_loss_bwd_dev(
            o["point_cloud_model"].data_ptr(), o["point_cloud_observed"].data_ptr(),
            o["point_cloud_weights"].data_ptr(), loop.rot_est.data_ptr(), loop.trans_est.data_ptr(),
            loop.pose[1 - loop.cur].data_ptr(), _lib.f3((0, 0, 0)), _lib.f3((1, 1, 1)),
            stp.rot_coord, stp.loss_type, stp.norm_term, 1.0, stp.grad_scale, gr.data_ptr(),
            gt.data_ptr(), B, stp.N, _lib.stream_ptr(None))
        assert rc == 0
        torch.cuda.synchronize()
        return gr, gt

    if sh["backward"]:
        gr, gt = direct_grads()
        assert torch.equal(gr, loop.step.grad_rot) and torch.equal(gt, loop.step.grad_trans), "eager"
    # ---- the same loop as a CUDA-graph replay: bit-identical final state
    eager = {"pose": loop.pose[loop.cur].clone(), "depth": loop.depth_rendered.clone(),
             "image": loop.image_rendered.clone(), "rot_est": loop.rot_est.clone(),
             "trans_est": loop.trans_est.clone(), "pose_prev": loop.pose[1 - loop.cur].clone()}
    for k, v in loop.step.outputs().items():
        if v is not None:
            eager[k] = v.clone()
    # the snapshot's clones run on the default stream; the stream below does not wait for it by
    # itself, and the captured loop's first launches overwrite what is still being cloned
    torch.cuda.synchronize()
    s = torch.cuda.Stream(device=cuda_device)
    with torch.cuda.stream(s):
        loop.capture(None, iters)
        for rep in range(2):
            loop.replay()
            s.synchronize()
            if sh["backward"]:
                gr, gt = direct_grads()
                assert torch.equal(gr, loop.step.grad_rot) and torch.equal(gt, loop.step.grad_trans), rep
            got = {"pose": loop.pose[loop.cur], "depth": loop.depth_rendered, "image": loop.image_rendered,
                   "rot_est": loop.rot_est, "trans_est": loop.trans_est,
                   "pose_prev": loop.pose[1 - loop.cur]}
            got.update({k: v for k, v in loop.step.outputs().items() if v is not None})
            for k in eager:
                x, y = eager[k], got[k]
                if x.dtype == torch.float32:
                    x, y = x.view(torch.int32), y.view(torch.int32)
                if not torch.equal(x, y):
                    bad = (x != y).reshape(x.shape[0], -1).any(dim=1)
                    rows = bad.nonzero().flatten()
                    raise AssertionError((cfg, rep, k, "pairs that differ: %d of %d, first %s"
                                          % (int(bad.sum()), x.shape[0], rows[:8].tolist()),
                                          eager[k][rows[:2]].tolist(), got[k][rows[:2]].tolist()))
    loop.check_status()
