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


def test_prep_image_and_depth(cuda_device):
    """image.transform and the depth decode of the loaders (lib/utils/image.py)."""
    import torch

    from mx_deepim_b200 import prep, synth

    rng = np.random.default_rng(5)
    B, H, W = 3, 480, 640
    im = rng.integers(0, 256, size=(B, H, W, 3), dtype=np.uint8)
    means = np.array([102.9801, 115.9465, 122.7717])  # B, G, R (config.network.PIXEL_MEANS)
    out = prep.transform_batch(torch.from_numpy(im).to(cuda_device), means)
    torch.cuda.synchronize()
    ref = np.zeros((B, 3, H, W))
    for b in range(B):  # image.py:717-720
        for i in range(3):
            ref[b, i] = im[b, :, :, 2 - i] - means[2 - i]
    assert np.array_equal(out.cpu().numpy(), ref.astype(np.float32))
    one = prep.transform(im[0], means)
    assert one.shape == (1, 3, H, W) and np.array_equal(one[0], ref[0].astype(np.float32))

    raw = rng.integers(0, 3000, size=(B, H, W)).astype(np.uint16)
    mask = rng.integers(0, 4, size=(B, H, W)).astype(np.uint8)
    idx = np.array([1, 2, 3], np.int32)
    t = torch.from_numpy(raw.view(np.int16)).to(cuda_device)
    d = prep.decode_depth(t, 1000.0)
    dm = prep.decode_depth(t, 1000.0, torch.from_numpy(mask).to(cuda_device),
                           torch.from_numpy(idx).to(cuda_device))
    torch.cuda.synchronize()
    ref_d = raw.astype(np.float32) / 1000.0  # image.py:200-206
    assert np.array_equal(d.cpu().numpy()[:, 0], ref_d)
    ref_m = raw.astype(np.float32)
    ref_m *= mask == idx.reshape(B, 1, 1)    # image.py:244
    assert np.array_equal(dm.cpu().numpy()[:, 0], ref_m / 1000.0)
    mk = prep.decode_mask(torch.from_numpy(mask).to(cuda_device), torch.from_numpy(idx).to(cuda_device))
    torch.cuda.synchronize()
    ref_k = np.zeros(mask.shape)
    ref_k[mask == idx.reshape(B, 1, 1)] = 1.0  # image.py:302-304
    assert np.array_equal(mk.cpu().numpy()[:, 0], ref_k.astype(np.float32))


def test_point_cloud_model_and_observed(cuda_device):
    import torch

    from mx_deepim_b200 import prep, synth

    rng = np.random.default_rng(9)
    pts_obj = rng.normal(size=(1234, 3)) * 0.05
    np.random.seed(4)
    ps, pw = prep.get_point_cloud_model(pts_obj, 3000)
    assert ps.shape == (1, 3, 3000) and pw.shape == (1, 3, 3000)
    assert pw[0, :, :1234].min() == 1 and pw[0, :, 1234:].max() == 0 and np.all(ps[0, :, 1234:] == 0)
    assert sorted(map(tuple, ps[0, :, :1234].T.tolist())) == sorted(map(tuple, pts_obj.tolist()))
    B = 4
    pts = np.repeat(ps.astype(np.float32), B, axis=0)
    poses = np.stack([synth.random_pose_pair(rng, synth.K_LINEMOD)[1] for _ in range(B)])
    out = prep.get_point_cloud_observed(torch.from_numpy(pts).to(cuda_device),
                                        torch.from_numpy(poses).to(cuda_device))
    torch.cuda.synchronize()
    for b in range(B):  # image.py:592-596
        R, T = poses[b][:, :3], poses[b][:, 3]
        ref = np.dot(R.astype(np.float64), pts[b].astype(np.float64)) + T.reshape((3, 1))
        np.testing.assert_allclose(out[b].cpu().numpy(), ref, rtol=0, atol=1e-6)


def test_calc_flow_device_matches_reference_golden(cuda_device):
    """prep.calc_flow (device, float64 numpy rule of flow.py:12-81) against the outputs of the
    REFERENCE's own calc_flow stored by tests/golden/make_golden.py, and against the oracle
    restatement on the standard_rep variant."""
    import os

    import helpers
    from mx_deepim_b200 import prep
    from oracle import flow_np

    gold = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    for seed in (0, 4):
        inp = helpers.make_flow_inputs(1, seed, occlude=(seed == 4))
        d_src = inp["depth_src"][0, 0].astype(np.float64)
        d_tgt = inp["depth_tgt"][0, 0].astype(np.float64)
        g = np.load(os.path.join(gold, "flow_calc_flow_seed%d.npz" % seed))
        flow, vis, xv = prep.calc_flow(d_src, inp["pose_src"][0], inp["pose_tgt"][0], helpers.K_LM, d_tgt)
        assert flow.shape == d_src.shape + (2,) and vis.shape == d_src.shape
        idx = np.flatnonzero(vis.reshape(-1))
        assert np.array_equal(idx, g["visible_idx"])          # visibility mask: bit-exact
        fa = flow.reshape(-1, 2)[idx]
        assert np.abs(fa - g["flow_at_visible"]).max() < 1e-4  # float32 store of the float64 value
        assert float(np.abs(flow.reshape(-1, 2)[np.setdiff1d(np.arange(vis.size), idx)]).max()) == 0
        assert xv.shape == (3, int(g["n_X_valid"]))
        np.testing.assert_allclose(xv.sum(axis=1), g["X_valid_sum"], rtol=1e-5)
        # standard representation (dw, dh) and the point list, vs the restatement
        f2, v2, x2 = prep.calc_flow(d_src, inp["pose_src"][0], inp["pose_tgt"][0], helpers.K_LM, d_tgt,
                                    standard_rep=True)
        rf, rv, rx = flow_np.calc_flow(d_src, inp["pose_src"][0], inp["pose_tgt"][0], helpers.K_LM,
                                       d_tgt, standard_rep=True)
        assert np.array_equal(v2, rv)
        assert np.abs(f2 - rf).max() < 1e-4
        np.testing.assert_allclose(x2, rx, rtol=1e-6, atol=1e-7)


def test_entry_points_refuse_strided_and_mistyped_tensors(cuda_device):
    """Tensors cross the C ABI as bare pointers, so the host mirrors must refuse views whose
    strides the kernels would misread (transposed loader output, channel slices) and wrong
    dtypes, instead of computing garbage silently."""
    import torch

    from mx_deepim_b200 import evaluation, flow, prep
    from mx_deepim_b200.render import DeviceRenderer
    from mx_deepim_b200 import synth

    dev = cuda_device
    pts_t = torch.zeros((2, 3000, 3), device=dev).transpose(1, 2)  # [B,3,N] view, not contiguous
    pose = torch.zeros((2, 3, 4), device=dev)
    with pytest.raises(ValueError):
        prep.get_point_cloud_observed(pts_t, pose)
    with pytest.raises(ValueError):
        prep.get_point_cloud_observed(pts_t.contiguous(), pose[:1])
    with pytest.raises(TypeError):
        prep.get_point_cloud_observed(pts_t.contiguous().double(), pose)
    depth4 = torch.zeros((2, 2, 480, 640), device=dev)
    with pytest.raises(ValueError):  # channel slice of a 2-channel tensor
        prep.calc_flow_batch(depth4[:, :1], depth4[:, 1:], torch.zeros((2, 3, 4), dtype=torch.float64, device=dev),
                             synth.K_LINEMOD)
    lab = torch.zeros((2, 480, 640), dtype=torch.uint8, device=dev)
    with pytest.raises(TypeError):
        prep.decode_mask(lab, torch.ones(2, dtype=torch.int64, device=dev))
    with pytest.raises(ValueError):
        evaluation.visib_mask(depth4[:, 0], depth4[:, 1], 0.015)
    with pytest.raises(ValueError):
        evaluation.pose_errors(pose, pose, pts_t, synth.K_LINEMOD)
    with pytest.raises(ValueError):
        flow.gpu_flow_dev(depth4[:, :1], depth4[:, 1:], pose, torch.eye(3, device=dev))
    ren = DeviceRenderer([synth.make_mesh(np.random.default_rng(0), level=1, texture_size=8)],
                         synth.K_LINEMOD, max_batch=2, device=dev)
    with pytest.raises(TypeError):
        ren.render_batch(torch.zeros(2, dtype=torch.int64, device=dev), pose)
    with pytest.raises(ValueError):
        ren.render_batch(torch.zeros(2, dtype=torch.int32, device=dev),
                         torch.zeros((2, 4, 3), device=dev).transpose(1, 2))
