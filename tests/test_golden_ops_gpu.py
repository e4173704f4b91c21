"""The CUDA plugins against golden vectors produced by the REFERENCE'S OWN operator files
(tests/golden/ops_reference.npz, made by tests/golden/make_golden_ops.py: deepim/operator_py/*.py
run unmodified under a numpy `mxnet` stand-in whose two MXNet built-ins are the sampler
restatement).  No oracle call in this file: the committed reference outputs are the checker.

Bars (north_star): zoom factors bit-exact -- wx, wy always; tx, ty wherever the golden run's
numpy-2 (NEP 50) scalar promotion coincides with the numpy-1.x promotion the product follows
(recorded per value in the fixture), else within 1.5 ulp(1); resampled values within 1e-4 abs;
0/1 masks equal on >= 99.999 % of the pixels; Transform3D within 1e-5 m, gradients 2e-4 rel;
one refine iteration against the reference driver's own outputs at the rasteriser bars.
"""
import os

import numpy as np
import pytest

import helpers

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
STRIDE = 8
TOL = 1e-4
K_FLAT = helpers.K_LM.flatten()
PM_FLAT = helpers.PIXEL_MEANS.flatten()


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLD, "ops_reference.npz"))


def _custom(dev, **kw):
    import torch

    from mx_deepim_b200 import operator, operator_py  # noqa: F401

    kw = {k: (torch.from_numpy(np.ascontiguousarray(v)).to(dev)
              if isinstance(v, np.ndarray) and k not in ("K", "pixel_means", "T_means", "T_stds")
              else v) for k, v in kw.items()}
    h = operator.Custom(**kw)
    torch.cuda.synchronize()
    return h


def bits(gold, key, shape):
    n = int(np.prod(shape))
    return np.unpackbits(gold[key + "__bits"])[:n].reshape(shape).astype(np.float32)


def check_plane(gold, key, got, tol=TOL):
    """`tol` is a scalar or one value per sample of the batch."""
    tol = np.broadcast_to(np.asarray(tol, np.float64), got.shape[:1])
    ref = gold[key + "__sample"]
    err = np.abs(got[..., ::STRIDE, ::STRIDE] - ref).reshape(got.shape[0], -1).max(axis=1)
    assert (err <= tol).all(), (key, err)
    s = got.astype(np.float64).sum(axis=(-2, -1))
    serr = np.abs(s - gold[key + "__sum"]).reshape(got.shape[0], -1).max(axis=1)
    assert (serr <= tol * got.shape[-1] * got.shape[-2]).all(), (key, serr)


def check_mask(gold, key, got):
    assert (got != bits(gold, key, got.shape)).mean() < 1e-5, key


def check_zoom_factor(got, ref, legacy_equal):
    g, r = got.view(np.uint32), ref.view(np.uint32)
    assert np.array_equal(g[:, :2], r[:, :2])
    assert np.array_equal(g[legacy_equal], r[legacy_equal])
    assert np.abs(got.astype(np.float64) - ref.astype(np.float64)).max() <= 1.5 * 2.0 ** -23


@pytest.mark.parametrize("tag,batch,seed,empty", [("a", 2, 5, ()), ("b", 4, 6, (1,))])
def test_zoom_mask_and_image_vs_reference_golden(cuda_device, gold, tag, batch, seed, empty):
    d = helpers.make_zoom_inputs(batch, seed, empty_rendered=empty)
    eq = gold["zoom_mask_%s/legacy_equal" % tag]
    h = _custom(cuda_device, op_type="ZoomMask", mask_observed=d["mask_observed"],
                mask_gt_observed=d["mask_gt_observed"], mask_rendered=d["mask_rendered"],
                src_pose=d["src_pose"], K=K_FLAT, height=480, width=640)
    out = dict(zip(("zoom_mask_observed", "zoom_mask_gt_observed", "zoom_mask_rendered",
                    "zoom_factor"), [o.cpu().numpy() for o in h.outputs]))
    check_zoom_factor(out["zoom_factor"], gold["zoom_mask_%s/zoom_factor" % tag], eq)
    for k in ("zoom_mask_observed", "zoom_mask_gt_observed", "zoom_mask_rendered"):
        check_mask(gold, "zoom_mask_%s/%s" % (tag, k), out[k])
    h = _custom(cuda_device, op_type="ZoomImage", image_observed=d["image_observed"],
                image_rendered=d["image_rendered"], src_pose=d["src_pose"], K=K_FLAT, height=480,
                width=640, pixel_means=PM_FLAT)
    o = [x.cpu().numpy() for x in h.outputs]
    check_zoom_factor(o[2], gold["zoom_image_%s/zoom_factor" % tag], eq)
    # pairs whose tx, ty coincide with the golden run's: the 1e-4 bar.  Elsewhere tx / ty differ by
    # <= 1.5 * 2^-23, i.e. the taps move by <= 5.7e-5 px, and these test images are per-pixel
    # random noise (|gradient| <= 255 / px): <= 1.5e-2
    tol = np.where(eq.all(axis=1), TOL, 2e-2)
    check_plane(gold, "zoom_image_%s/zoom_image_observed" % tag, o[0], tol)
    check_plane(gold, "zoom_image_%s/zoom_image_rendered" % tag, o[1], tol)


def test_with_factor_family_vs_reference_golden(cuda_device, gold):
    d = helpers.make_zoom_inputs(2, 5)
    zf = gold["zoom_mask_a/zoom_factor"]  # fed as data: identical sampling coordinates
    for hl in (False, True):
        h = _custom(cuda_device, op_type="ZoomImageWithFactor", zoom_factor=zf,
                    image_observed=d["image_observed"], image_rendered=d["image_rendered"],
                    height=480, width=640, pixel_means=PM_FLAT, high_light_center=hl)
        check_plane(gold, "zoom_image_wf_hl%d/zoom_image_observed" % hl, h.outputs[0].cpu().numpy())
        check_plane(gold, "zoom_image_wf_hl%d/zoom_image_rendered" % hl, h.outputs[1].cpu().numpy())
    h = _custom(cuda_device, op_type="ZoomDepth", zoom_factor=zf,
                depth_observed=d["depth_observed"], depth_rendered=d["depth_rendered"])
    check_plane(gold, "zoom_depth/zoom_depth_observed", h.outputs[0].cpu().numpy())
    check_plane(gold, "zoom_depth/zoom_depth_rendered", h.outputs[1].cpu().numpy())
    h = _custom(cuda_device, op_type="ZoomFlow", zoom_factor=zf, flow=d["flow"],
                flow_weights=d["flow_weights"], b_inv_zoom=False)
    zoomed = h.outputs[0].cpu().numpy()
    check_plane(gold, "zoom_flow/zoom_flow", zoomed, TOL * 4)  # 1e-4 px scaled by 1 / wx <= 4
    check_mask(gold, "zoom_flow/zoom_flow_weights", h.outputs[1].cpu().numpy())
    h = _custom(cuda_device, op_type="ZoomMaskWithFactor", zoom_factor=zf,
                mask=d["depth_rendered"], b_inv_zoom=False)
    check_mask(gold, "zoom_mask_wf/zoom_mask", h.outputs[0].cpu().numpy())


def test_zoom_trans_vs_reference_golden(cuda_device, gold):
    import torch

    zf = gold["zoom_mask_a/zoom_factor"]
    td, og = gold["zoom_trans/trans_delta"], gold["zoom_trans/out_grad"]
    for inv in (False, True):
        for zg in (False, True):
            h = _custom(cuda_device, op_type="ZoomTrans", zoom_factor=zf, trans_delta=td,
                        b_inv_zoom=inv, b_zoom_grad=zg)
            assert np.array_equal(h.outputs[0].cpu().numpy().view(np.uint32),
                                  gold["zoom_trans/fwd_inv%d_zg%d" % (inv, zg)].view(np.uint32))
            gz, gt = h.backward([torch.from_numpy(og).to(cuda_device)])
            assert float(gz.abs().sum()) == 0.0
            assert np.array_equal(gt.cpu().numpy().view(np.uint32),
                                  gold["zoom_trans/bwd_inv%d_zg%d" % (inv, zg)].view(np.uint32))


@pytest.mark.parametrize("rot_coord", ["MODEL", "CAMERA", "CAMERA_NEW", "NAIVE"])
def test_transform3d_vs_reference_golden(cuda_device, gold, rot_coord):
    import torch

    p = helpers.make_points_inputs(4, 11, num_points=3000, num_keep=2500)
    og = np.random.default_rng(78).normal(0, 1.0, size=p["points"].shape).astype(np.float32)
    h = _custom(cuda_device, op_type="Transform3D", point_cloud=p["points"],
                rotation=p["rotation"], translation=p["translation"], pose_src=p["pose_src"],
                T_means=np.zeros(3, np.float32), T_stds=np.ones(3, np.float32),
                rot_coord=rot_coord)
    ref = gold["transform3d/%s/fwd" % rot_coord]
    assert np.abs(h.outputs[0].cpu().numpy() - ref).max() < 1e-5
    grads = h.backward([torch.from_numpy(og).to(cuda_device)])
    torch.cuda.synchronize()
    assert float(grads[0].abs().sum()) == 0.0 and float(grads[3].abs().sum()) == 0.0
    for got, key in ((grads[1], "grad_rotation"), (grads[2], "grad_translation")):
        r = gold["transform3d/%s/%s" % (rot_coord, key)]
        assert np.abs(got.cpu().numpy() - r).max() <= 2e-4 * max(1.0, np.abs(r).max()), key


def test_refine_iteration_vs_reference_driver_golden(cuda_device):
    """One iteration of the on-device loop (pose update -> raster -> post -> labels -> KT -> GT
    flow; RefineLoop.iterate) against what the reference's own batchUpdaterPyMulti.forward
    produced for the same batch and the same 'network predictions'
    (tests/golden/updater_reference.npz, see make_golden_updater.py for its stand-ins)."""
    import torch

    from mx_deepim_b200 import pipeline, synth
    from mx_deepim_b200 import pose as P

    g = np.load(os.path.join(GOLD, "updater_reference.npz"))
    inp = helpers.make_updater_inputs(3, 31)
    B = 3
    rot = torch.from_numpy(inp["rot_est"]).to(cuda_device)
    trans = torch.from_numpy(inp["trans_est"]).to(cuda_device)
    holder = {}

    def predict(it, rot_gt, trans_gt):
        holder["loop"].rot_est.copy_(rot)
        holder["loop"].trans_est.copy_(trans)

    loop = pipeline.RefineLoop(inp["meshes"], B, cuda_device, K=helpers.K_LM, predict=predict)
    holder["loop"] = loop
    rng = np.random.default_rng(0)
    pts = np.zeros((B, 3, synth.NUM_3D_SAMPLE), np.float32)
    wts = np.zeros_like(pts)
    for b in range(B):
        pts[b], wts[b] = synth.mesh_points(inp["meshes"][inp["class_index"][b]], rng)
    mask_gt = (inp["depth_gt_observed"] > 0).astype(np.float32)
    host = dict(class_index=inp["class_index"], src_pose=inp["src_pose"], tgt_pose=inp["tgt_pose"],
                image_observed=np.zeros((B, 3, 480, 640), np.float32),
                depth_gt_observed=inp["depth_gt_observed"], mask_gt_observed=mask_gt,
                mask_observed=mask_gt.copy(), point_cloud_model=pts, point_cloud_weights=wts,
                point_cloud_observed=pts.copy(), Kinv=P.make_Kinv(helpers.K_LM))
    loop.set_batch(host)
    loop.iterate(0)
    torch.cuda.synchronize()
    np.testing.assert_allclose(loop.pose[loop.cur].cpu().numpy(), g["src_pose"], rtol=2e-6, atol=3e-7)
    np.testing.assert_allclose(loop.rot_gt.cpu().numpy(), g["rot"], atol=3e-6)
    np.testing.assert_allclose(loop.trans_gt.cpu().numpy(), g["trans"], rtol=1e-5, atol=2e-6)
    depth = loop.depth_rendered.cpu().numpy()
    n = depth.size
    fg = np.unpackbits(g["foreground__bits"])[:n].reshape(depth.shape).astype(bool)
    assert ((depth > 0) == fg).mean() >= 0.999  # north_star: >= 99.9 % mask agreement
    mask = np.unpackbits(g["mask_rendered__bits"])[:n].reshape(depth.shape)
    assert (loop.mask_rendered.cpu().numpy() == mask).mean() >= 0.999
    S = 4
    d, dg = depth[..., ::S, ::S], g["depth_rendered__sample"]
    both = (d > 0) & (dg > 0)
    assert both.sum() > 200
    assert ((np.abs(d[both] - dg[both]) / dg[both]) <= 1e-5).mean() >= 0.999  # 1e-5 relative
    im, img = loop.image_rendered.cpu().numpy()[..., ::S, ::S], g["image_rendered__sample"]
    b3 = np.broadcast_to(both, im.shape)
    assert (np.abs(im[b3] - img[b3]) < 1e-3).mean() >= 0.995
    bg = np.broadcast_to(~(d > 0) & ~(dg > 0), im.shape)
    assert np.array_equal(im[bg], img[bg])
    w = np.unpackbits(g["flow_weights__bits"])[:2 * n].reshape((B, 2, 480, 640))
    valid = loop.step.valid.cpu().numpy()
    assert (valid == w[:, :1]).mean() >= 0.999
    fl, flg = loop.step.flow.cpu().numpy()[..., ::S, ::S], g["flow__sample"]
    v = np.broadcast_to((w[:, :1, ::S, ::S] > 0) & (valid[..., ::S, ::S] > 0), fl.shape)
    assert v.sum() > 100 and np.abs(fl[v] - flg[v]).max() < 2e-3


def test_loader_and_test_update_vs_reference_golden(cuda_device):
    """Device image transform, observed point cloud and the test-time mask policies against the
    outputs of the reference's own lib/utils/image.py and data_pair.update_data_batch
    (tests/golden/prep_reference.npz)."""
    import torch

    from mx_deepim_b200 import pipeline, prep

    g = np.load(os.path.join(GOLD, "prep_reference.npz"))
    inp = helpers.make_prep_inputs(21)
    t = prep.transform(inp["bgr"], inp["pixel_means"])
    assert t.shape == (1, 3, 480, 640)
    assert np.array_equal(t[..., ::4, ::4].astype(np.float32), g["transform"])
    # the product stores float32 (the batch blob dtype); the reference sums its float64 temporaries
    np.testing.assert_allclose(t.astype(np.float64).sum(axis=(-2, -1)), g["transform_sum"], rtol=1e-6)
    # update_data_batch's image blob = resize (identity at 480x640) + transform of a float frame
    assert np.array_equal(t[..., ::4, ::4].astype(np.float32), g["update/image_rendered"])
    # the stored array is the loader's TRANSPOSED view (image.py:586): a non-contiguous tensor
    # must be refused, not read with the wrong strides
    strided = torch.from_numpy(g["points_sample"].astype(np.float32)).to(cuda_device)
    pose = torch.from_numpy(inp["pose"][None]).to(cuda_device)
    if not strided.is_contiguous():
        with pytest.raises(ValueError):
            prep.get_point_cloud_observed(strided, pose)
    pts = torch.from_numpy(np.ascontiguousarray(g["points_sample"], np.float32)).to(cuda_device)
    obs = prep.get_point_cloud_observed(pts, pose)
    torch.cuda.synchronize()
    np.testing.assert_allclose(obs[0].cpu().numpy(), g["points_observed"], rtol=0, atol=1e-6)
    n = 480 * 640
    for policy, src in (("box_rendered", "mask_rendered"), ("box_observed", "mask_observed")):
        ref = np.unpackbits(g["update_%s/mask_observed__bits" % policy])[:n].reshape(480, 640)
        m = torch.from_numpy(inp[src][None, None]).to(cuda_device)
        out, _ = pipeline.mask_box(m)
        torch.cuda.synchronize()
        assert np.array_equal(out[0, 0].cpu().numpy(), ref.astype(np.float32)), policy


def test_dataset_files_to_device_blobs_vs_reference_loaders(cuda_device, tmp_path):
    """Files in the dataset's on-disk formats -> dataio.read_pair / stack_pairs (storage dtypes,
    pinned) -> upload -> device decode, against the float blobs the reference's own loaders made
    from the same files (tests/golden/dataio_reference.npz): bit-exact."""
    import torch

    from mx_deepim_b200 import dataio

    g = np.load(os.path.join(GOLD, "dataio_reference.npz"))
    pairdb = helpers.write_pair_files(tmp_path, 41)
    batch = dataio.stack_pairs([dataio.read_pair(r) for r in pairdb], pin=True)
    out = dataio.decode_batch(batch, cuda_device, depth_factor=1000.0,
                              pixel_means_bgr=[102.9801, 115.9465, 122.7717], mask_inputs=True)
    torch.cuda.synchronize()
    S = 4
    for k in ("image_observed", "image_rendered", "depth_gt_observed", "depth_observed",
              "depth_rendered"):
        got = out[k].cpu().numpy()
        assert got.dtype == np.float32 and got.shape[2:] == (480, 640), k
        assert np.array_equal(got[..., ::S, ::S], g[k]), k
    for k in ("depth_gt_observed", "depth_observed", "depth_rendered"):
        np.testing.assert_allclose(out[k].double().sum(dim=(-2, -1)).cpu().numpy(), g[k + "_sum"],
                                   rtol=1e-12)
    m = out["mask_gt_observed"].cpu().numpy()
    assert m[0].sum() == 200 * 250 and m[1].sum() == 190 * 270  # label == mask_idx, per pair
    np.testing.assert_array_equal(out["pose_observed"].cpu().numpy(), g["poses"].astype(np.float32))


@pytest.mark.parametrize("mode", ["MODEL", "CAMERA", "CAMERA_NEW"])
def test_pose_update_modes_vs_reference_golden(cuda_device, mode):
    """rc_pose_update_dev for every rot_coord mode and non-trivial T_means / T_stds against the
    reference's own RT_transform / calc_RT_delta / calc_se3 outputs
    (tests/golden/pose_modes_reference.npz)."""
    import torch

    from mx_deepim_b200 import pipeline

    g = np.load(os.path.join(GOLD, "pose_modes_reference.npz"))
    inp = helpers.make_pose_mode_inputs(51)
    t = lambda a: torch.from_numpy(a).to(cuda_device)  # noqa: E731
    out = pipeline.pose_update(t(inp["src_pose"]), t(inp["rot_est"]), t(inp["trans_est"]),
                               t(inp["tgt_pose"]), helpers.K_LM, T_means=(0.01, -0.02, 0.05),
                               T_stds=(0.5, 0.8, 1.5), rot_coord=mode)
    torch.cuda.synchronize()
    pose, rot, trans, KT = [o.cpu().numpy() for o in out]
    np.testing.assert_allclose(pose, g[mode + "/pose"], rtol=2e-6, atol=3e-7)
    np.testing.assert_allclose(rot, g[mode + "/rot"], rtol=0, atol=3e-6)
    np.testing.assert_allclose(trans, g[mode + "/trans"], rtol=1e-5, atol=3e-6)
    np.testing.assert_allclose(KT, g[mode + "/KT"], rtol=2e-6, atol=3e-4)


def test_mask_preparation_vs_reference_loader_golden(cuda_device, tmp_path):
    """Iteration-0 masks on the device (prep.get_pair_mask: label decode, bounding rectangle,
    mask_dilate with the reference's random draws, mask_rendered) against the reference's own
    get_pair_mask / mask_dilate outputs (tests/golden/masks_reference.npz): bit for bit."""
    import cv2
    import torch

    from mx_deepim_b200 import prep

    g = np.load(os.path.join(GOLD, "masks_reference.npz"))
    pairdb = helpers.write_pair_files(tmp_path, 41)
    label = np.stack([cv2.imread(r["mask_gt_observed"], cv2.IMREAD_UNCHANGED) for r in pairdb])
    raw = np.stack([cv2.imread(r["depth_rendered"], cv2.IMREAD_UNCHANGED) for r in pairdb])
    lab_d = torch.from_numpy(label).to(cuda_device)
    idx_d = torch.tensor([r["mask_idx"] for r in pairdb], dtype=torch.int32, device=cuda_device)
    depth = prep.decode_depth(torch.from_numpy(raw.view(np.int16)).to(cuda_device), 1000.0)
    n = 2 * 480 * 640
    for phase, init in (("train", "box_gt"), ("test", "box_rendered")):
        for dil in (False, True):
            tag = "%s_%s_dil%d" % (phase, init, dil)
            np.random.seed(77)
            obs, gt, rend = prep.get_pair_mask(lab_d, idx_d, depth, phase=phase, init_mask=init,
                                               dilate=dil)
            torch.cuda.synchronize()
            ref_o = np.unpackbits(g[tag + "/mask_observed"])[:n].reshape(2, 1, 480, 640)
            ref_g = np.unpackbits(g[tag + "/mask_gt_observed"])[:n].reshape(2, 1, 480, 640)
            assert np.array_equal(obs.cpu().numpy(), ref_o.astype(np.float32)), tag
            assert np.array_equal(gt.cpu().numpy(), ref_g.astype(np.float32)), tag
            r = rend.cpu().numpy()
            assert np.array_equal(r[..., ::4, ::4], g[tag + "/mask_rendered"]), tag
            np.testing.assert_allclose(r.astype(np.float64).sum(axis=(-2, -1)),
                                       g[tag + "/mask_rendered_sum"], rtol=1e-12)
    rng = np.random.default_rng(9)
    masks = (rng.random((12, 96, 128)) > 0.97).astype(np.float64)
    masks[3] *= 0.5
    masks[4, 10:20, 10:20] = 3.0
    np.random.seed(5)
    draws = [prep.draw_dilate(10) for _ in range(12)]
    out = prep.mask_dilate(torch.from_numpy(masks[:, None].astype(np.float32)).to(cuda_device), draws)
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy()[:, 0], g["dilate/out"].astype(np.float32))
    with pytest.raises(Exception):
        prep.get_pair_mask(lab_d, idx_d, depth, phase="train", init_mask="mask_gt")
