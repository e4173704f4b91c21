"""The oracle's plugin restatements against golden vectors produced by the REFERENCE'S OWN
operator files (deepim/operator_py/*.py run unmodified under a numpy-backed `mxnet` stand-in,
tests/golden/make_golden_ops.py + mx_stub.py).  What this pins: every line of Python / numpy the
plugins execute (crop box, zoom factor, thresholds, mean handling, centre spot, inverse-zoom
affine, flow rescale, weight binarisation, ZoomTrans, Transform3D forward and gradients) and
the Prop metadata.  What it does not pin: MXNet's GridGenerator / BilinearSampler, which the
stand-in serves from the same restatement ("parity unpinned").

The golden run used numpy 2.x, so mixed np.float32 / Python-scalar expressions carry NEP-50
values; `zoom_np.promotion("nep50")` reproduces them bit-exactly, the default "legacy" mode
(numpy 1.x, what the product implements) differs by at most 1.8e-7 absolute on tx / ty.
"""
import json
import os

import numpy as np
import pytest

import helpers
from oracle import transform3d_np, zoom_np

GOLD = os.path.join(os.path.dirname(__file__), "golden")
STRIDE = 8


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLD, "ops_reference.npz"))


def bits(gold, key, shape):
    n = int(np.prod(shape))
    return np.unpackbits(gold[key + "__bits"])[:n].reshape(shape).astype(np.float32)


def same_bits(a, b):
    return np.array_equal(np.ascontiguousarray(a, np.float32).view(np.uint32),
                          np.ascontiguousarray(b, np.float32).view(np.uint32))


def check_plane(gold, key, got):
    assert same_bits(got[..., ::STRIDE, ::STRIDE], gold[key + "__sample"]), key
    np.testing.assert_allclose(got.astype(np.float64).sum(axis=(-2, -1)), gold[key + "__sum"],
                               rtol=1e-12, err_msg=key)


@pytest.mark.parametrize("tag,batch,seed,empty", [("a", 2, 5, ()), ("b", 4, 6, (1,))])
def test_zoom_mask_and_image_vs_reference_plugins(gold, tag, batch, seed, empty):
    d = helpers.make_zoom_inputs(batch, seed, empty_rendered=empty)
    with zoom_np.promotion("nep50"):
        r0, r1, r2, zf, _ = zoom_np.zoom_mask_forward(
            d["mask_observed"], d["mask_gt_observed"], d["mask_rendered"], d["src_pose"],
            helpers.K_LM)
        ia, ir, zfi, _ = zoom_np.zoom_image_forward(
            d["image_observed"], d["image_rendered"], d["src_pose"], helpers.K_LM,
            helpers.PIXEL_MEANS[::-1])
    ref_zf = gold["zoom_mask_%s/zoom_factor" % tag]
    assert same_bits(zf, ref_zf)
    shape = d["mask_observed"].shape
    for name, got in (("zoom_mask_observed", r0), ("zoom_mask_gt_observed", r1),
                      ("zoom_mask_rendered", r2)):
        assert np.array_equal(got, bits(gold, "zoom_mask_%s/%s" % (tag, name), shape)), name
    assert not gold["zoom_mask_%s/grad_abs_sum" % tag].any()  # zoom_mask.py:136-140
    assert same_bits(zfi, gold["zoom_image_%s/zoom_factor" % tag])
    check_plane(gold, "zoom_image_%s/zoom_image_observed" % tag, ia)
    check_plane(gold, "zoom_image_%s/zoom_image_rendered" % tag, ir)

    # numpy-1.x promotion (the product's): wx, wy identical; tx, ty flagged where they differ
    legacy = zoom_np.zoom_mask_forward(d["mask_observed"], d["mask_gt_observed"],
                                       d["mask_rendered"], d["src_pose"], helpers.K_LM)[3]
    eq = legacy.view(np.uint32) == ref_zf.view(np.uint32)
    assert np.array_equal(eq, gold["zoom_mask_%s/legacy_equal" % tag])
    assert eq[:, :2].all()
    # three float32 roundings of a value in [0, 2) vs one: <= 1.5 ulp(1) absolute on x * 2 - 1
    assert np.abs(legacy.astype(np.float64) - ref_zf.astype(np.float64)).max() <= 1.5 * 2.0 ** -23
    for b in empty:  # no rendered pixel: centre from the observed box, float64 in both modes
        assert eq[b].all()


def test_with_factor_family_vs_reference_plugins(gold):
    d = helpers.make_zoom_inputs(2, 5)
    zf = gold["zoom_mask_a/zoom_factor"]
    pm = helpers.PIXEL_MEANS[::-1]
    with zoom_np.promotion("nep50"):
        for hl in (False, True):
            a, r = zoom_np.zoom_image_with_factor_forward(zf, d["image_observed"],
                                                          d["image_rendered"], pm, hl)
            check_plane(gold, "zoom_image_wf_hl%d/zoom_image_observed" % hl, a)
            check_plane(gold, "zoom_image_wf_hl%d/zoom_image_rendered" % hl, r)
        do, dr = zoom_np.zoom_depth_forward(zf, d["depth_observed"], d["depth_rendered"])
        check_plane(gold, "zoom_depth/zoom_depth_observed", do)
        check_plane(gold, "zoom_depth/zoom_depth_rendered", dr)
        fl, w = zoom_np.zoom_flow_forward(zf, d["flow"], d["flow_weights"])
        check_plane(gold, "zoom_flow/zoom_flow", fl)
        assert np.array_equal(w, bits(gold, "zoom_flow/zoom_flow_weights", w.shape))
        back = zoom_np.zoom_flow_forward(zf, fl, b_inv_zoom=True)
        check_plane(gold, "zoom_flow_inv/zoom_flow", back)
        m = zoom_np.zoom_mask_with_factor_forward(zf, d["depth_rendered"])
        assert np.array_equal(m, bits(gold, "zoom_mask_wf/zoom_mask", m.shape))
        mi = zoom_np.zoom_mask_with_factor_forward(zf, m, b_inv_zoom=True)
        assert np.array_equal(mi, bits(gold, "zoom_mask_wf_inv/zoom_mask", mi.shape))
    # legacy promotion of the inverse affine: same taps except at coordinates within float32
    # rounding of a pixel boundary
    back_legacy = zoom_np.zoom_flow_forward(zf, fl, b_inv_zoom=True)
    assert np.abs(back_legacy - back).max() < 1e-3
    mi_legacy = zoom_np.zoom_mask_with_factor_forward(zf, m, b_inv_zoom=True)
    assert (mi_legacy != mi).mean() < 1e-4


def test_zoom_trans_vs_reference_plugin(gold):
    zf = gold["zoom_mask_a/zoom_factor"]
    td, og = gold["zoom_trans/trans_delta"], gold["zoom_trans/out_grad"]
    for inv in (False, True):
        for zg in (False, True):
            assert same_bits(zoom_np.zoom_trans_forward(zf, td, inv),
                             gold["zoom_trans/fwd_inv%d_zg%d" % (inv, zg)])
            assert same_bits(zoom_np.zoom_trans_backward(zf, og, inv, zg),
                             gold["zoom_trans/bwd_inv%d_zg%d" % (inv, zg)])
    assert same_bits(gold["zoom_trans/fwd_req_add_from_zero"], gold["zoom_trans/fwd_inv0_zg0"])


@pytest.mark.parametrize("rot_coord", ["MODEL", "CAMERA", "CAMERA_NEW", "NAIVE"])
def test_transform3d_vs_reference_plugin(gold, rot_coord):
    p = helpers.make_points_inputs(4, 11, num_points=3000, num_keep=2500)
    og = np.random.default_rng(78).normal(0, 1.0, size=p["points"].shape).astype(np.float32)
    tm, ts = np.zeros(3, np.float32), np.ones(3, np.float32)
    fwd = transform3d_np.forward(p["points"], p["rotation"], p["translation"], p["pose_src"],
                                 tm, ts, rot_coord)
    ref = gold["transform3d/%s/fwd" % rot_coord]
    assert np.abs(fwd - ref).max() < 1e-5  # metres; float32 batch_dot association differs
    g_rot, g_tr = transform3d_np.backward(og, p["points"], p["rotation"], p["translation"],
                                          p["pose_src"], tm, ts, rot_coord)[:2]
    for got, key in ((g_rot, "grad_rotation"), (g_tr, "grad_translation")):
        r = gold["transform3d/%s/%s" % (rot_coord, key)]
        assert np.abs(got - r).max() <= 2e-4 * max(1.0, np.abs(r).max()), key


def test_prop_metadata_vs_reference_plugins():
    """list_arguments / list_outputs / infer_shape of the product's Props (host logic, no GPU)
    against what the reference's Props returned in the golden run."""
    from mx_deepim_b200 import operator, operator_py  # noqa: F401

    with open(os.path.join(GOLD, "ops_reference.json")) as f:
        meta = json.load(f)
    k_str, pm_str = str(helpers.K_LM.flatten()), str(helpers.PIXEL_MEANS.flatten())
    kwargs = {
        "ZoomMask": dict(K=k_str, height="480", width="640"),
        "ZoomImage": dict(K=k_str, height="480", width="640", pixel_means=pm_str),
        "ZoomImageWithFactor": dict(height="480", width="640", pixel_means=pm_str,
                                    high_light_center="True"),
        "ZoomDepth": dict(height="480", width="640"),
        "ZoomFlow": dict(height="480", width="640", b_inv_zoom="False"),
        "ZoomFlow_inv": dict(height="480", width="640", b_inv_zoom="True"),
        "ZoomMaskWithFactor": dict(height="480", width="640", b_inv_zoom="False"),
        "ZoomTrans": dict(b_inv_zoom="True", b_zoom_grad="True"),
        "Transform3D": dict(T_means="[0. 0. 0.]", T_stds="[1. 1. 1.]", rot_coord="NAIVE"),
    }
    import inspect

    for name, ref in meta.pop("__prop_init__").items():  # kwargs names + defaults of each Prop
        params = list(inspect.signature(operator.get_prop_class(name).__init__).parameters.values())
        got = [[q.name, None if q.default is inspect.Parameter.empty else repr(q.default)]
               for q in params[1:]]
        assert got == ref, (name, got, ref)
    assert set(kwargs) == set(meta)
    for name, kw in kwargs.items():
        prop = operator.get_prop_class(name.split("_")[0])(**kw)
        m = meta[name]
        assert prop.list_arguments() == m["arguments"], name
        assert prop.list_outputs() == m["outputs"], name
        ins, outs, aux = prop.infer_shape([list(s) for s in m["in_shapes"]])
        assert [list(s) for s in ins] == m["in_shapes"], name
        assert [list(s) for s in outs] == m["out_shapes"], name
        assert list(aux) == []


# ------------------------------------------------------------------ renderer boundary
def test_renderer_matrices_vs_reference_modules():
    """oracle/raster_np.py's lifted my_compute_calib_proj / _get_view_mtx against what every
    reference renderer class returned (tests/golden/make_golden_render.py)."""
    from oracle import raster_np

    g = np.load(os.path.join(GOLD, "render_reference.npz"))
    with open(os.path.join(GOLD, "render_reference.json")) as f:
        meta = json.load(f)
    Ks = {"lm": helpers.K_LM, "other": np.array([[1066.778, 0, 312.9869], [0, 1067.487, 241.3109],
                                                  [0, 0, 1]], np.float32)}
    poses = g["poses"]
    for name in meta:
        for kn, K in Ks.items():
            for (w, h) in ((640, 480), (320, 200)):
                ref = g["%s/proj_%s_float64_%dx%d" % (name, kn, w, h)]
                got = raster_np.my_compute_calib_proj(K.astype(np.float64), w, h, 0.25, 6.0)
                assert np.array_equal(got, ref), (name, kn, w, h)
                # a float32 K under numpy 2 (NEP 50) rounds the frustum to float32: same matrix
                # to float32 precision, which is all the GL uniform keeps
                ref32 = g["%s/proj_%s_float32_%dx%d" % (name, kn, w, h)]
                np.testing.assert_allclose(got, ref32, rtol=3e-6, atol=1e-7)
        views = np.asarray([raster_np._get_view_mtx(p[:, :3], p[:, 3]) for p in poses])
        assert views.dtype == np.float32 and np.array_equal(views, g[name + "/view"]), name


def test_renderer_signatures_vs_reference_modules():
    """Constructor / render() parameter names and defaults of the drop-in renderer classes
    (mx_deepim_b200/render_glumpy/<module>.<Class>) against the reference's: the
    reference's parameters must come first, in order, with equal defaults; the product may
    append keyword-only extras (mesh=, max_batch=, device=)."""
    import importlib
    import inspect

    with open(os.path.join(GOLD, "render_reference.json")) as f:
        meta = json.load(f)
    for module, m in meta.items():
        mod = importlib.import_module("mx_deepim_b200.render_glumpy." + module)
        cls = getattr(mod, m["class"])
        for fn_name, key in (("__init__", "init"), ("render", "render")):
            params = list(inspect.signature(getattr(cls, fn_name)).parameters.values())[1:]
            got = [[p.name, None if p.default is inspect.Parameter.empty else repr(p.default)]
                   for p in params]
            ref = m[key]
            assert got[:len(ref)] == ref, (module, fn_name, got, ref)
            assert all(d is not None for _, d in got[len(ref):]), (module, fn_name)


# ------------------------------------------------------------------ batch updater (a22)
def test_updater_restatement_vs_reference_driver():
    """oracle/updater_np.updater_forward against the reference's own batchUpdaterPyMulti.forward
    (tests/golden/make_golden_updater.py: the driver run unmodified with stand-ins for mxnet,
    gpu_flow and the GL renderer).  The golden run's numpy 2 keeps `2.0 / Nq` of quat2mat in
    float32 where the restatement follows numpy 1.x (float64): poses agree to float32 rounding,
    so frames are compared at the north_star bars rather than bit for bit."""
    from oracle import updater_np

    g = np.load(os.path.join(GOLD, "updater_reference.npz"))
    inp = helpers.make_updater_inputs(3, 31)
    pm = helpers.PIXEL_MEANS[[2, 1, 0]]
    ref = updater_np.updater_forward(inp["meshes"], inp["class_index"], inp["src_pose"],
                                     inp["tgt_pose"], inp["rot_est"], inp["trans_est"],
                                     inp["depth_gt_observed"], helpers.K_LM, pm, np.zeros(3),
                                     np.ones(3))
    assert list(g["pred_names"]).index("rot_est") == 4  # get_names(): flow preds come first
    np.testing.assert_allclose(ref["src_pose"], g["src_pose"], rtol=0, atol=3e-7)
    np.testing.assert_allclose(ref["rot"], g["rot"], rtol=0, atol=2e-6)
    np.testing.assert_allclose(ref["trans"], g["trans"], rtol=1e-5, atol=2e-6)
    shape = ref["depth_rendered"].shape
    n = int(np.prod(shape))
    fg = np.unpackbits(g["foreground__bits"])[:n].reshape(shape).astype(bool)
    assert fg.sum() > 3000
    assert ((ref["depth_rendered"] > 0) == fg).mean() >= 0.9999
    mask = np.unpackbits(g["mask_rendered__bits"])[:n].reshape(shape)
    assert (ref["mask_rendered"] == mask).mean() >= 0.9999
    S = 4
    d, dg = ref["depth_rendered"][..., ::S, ::S], g["depth_rendered__sample"]
    both = (d > 0) & (dg > 0)
    assert (np.abs(d[both] - dg[both]) / dg[both]).max() <= 1e-5
    im, img = ref["image_rendered"][..., ::S, ::S], g["image_rendered__sample"]
    b3 = np.broadcast_to(both, im.shape)
    assert (np.abs(im[b3] - img[b3]) < 1e-3).mean() >= 0.995  # a texel boundary may flip
    bg = np.broadcast_to(~(d > 0) & ~(dg > 0), im.shape)
    assert np.array_equal(im[bg], img[bg])  # background = -pixel_means exactly
    w = np.unpackbits(g["flow_weights__bits"])[:2 * n].reshape(ref["flow_weights"].shape)
    assert np.array_equal(w[:, 0], w[:, 1])  # np.tile(valid, [1, 2, 1, 1])
    assert (ref["flow_weights"] == w).mean() >= 0.9999
    fl, flg = ref["flow"][..., ::S, ::S], g["flow__sample"]
    v = np.broadcast_to((w[:, :1, ::S, ::S] > 0) & (ref["flow_weights"][:, :1, ::S, ::S] > 0),
                        fl.shape)
    assert v.sum() > 100 and np.abs(fl[v] - flg[v]).max() < 2e-3


# ------------------------------------------------------------------ loaders / test-time update
def test_loader_and_test_update_vs_reference_code():
    """tests/golden/prep_reference.npz (lib/utils/image.py and lib/pair_matching/data_pair.py run
    unmodified): the oracle's lift of update_data_batch's mask policies and the product's host
    mirror of get_point_cloud_model (same np.random stream) are bit-identical to it."""
    from mx_deepim_b200 import prep
    from oracle import updater_np

    g = np.load(os.path.join(GOLD, "prep_reference.npz"))
    inp = helpers.make_prep_inputs(21)
    n = 480 * 640
    for policy in ("box_rendered", "box_observed", "init"):
        ref = np.unpackbits(g["update_%s/mask_observed__bits" % policy])[:n].reshape(480, 640)
        got = updater_np.update_mask_observed(policy, inp["mask_observed"], inp["mask_rendered"])
        assert np.array_equal(got, ref), policy
        rend = np.unpackbits(g["update_%s/mask_rendered__bits" % policy])[:n].reshape(480, 640)
        assert np.array_equal(rend, inp["mask_rendered"])
    assert np.array_equal(g["update/src_pose"][0], inp["pose"])
    assert np.array_equal(g["update/depth_rendered"][0, 0], inp["depth"][::4, ::4])
    np.random.seed(4)
    ps, pw = prep.get_point_cloud_model(inp["points_obj"], 3000)
    assert np.array_equal(ps, g["points_sample"]) and np.array_equal(pw, g["points_weights"])


@pytest.mark.parametrize("mode", ["MODEL", "CAMERA", "CAMERA_NEW"])
def test_pose_block_modes_vs_reference_rt_transform(mode):
    """oracle/updater_np.pose_update for every rot_coord mode and non-trivial T_means / T_stds
    against the reference's own RT_transform / calc_RT_delta / calc_se3
    (tests/golden/make_golden_pose_modes.py).  The golden run's numpy 2 keeps quat2mat's
    `2.0 / Nq` in float32 (numpy 1.x: float64), hence float32-rounding tolerances."""
    from oracle import updater_np

    g = np.load(os.path.join(GOLD, "pose_modes_reference.npz"))
    inp = helpers.make_pose_mode_inputs(51)
    means, stds = np.array([0.01, -0.02, 0.05]), np.array([0.5, 0.8, 1.5])
    pose, rot, trans, KT = updater_np.pose_update(inp["src_pose"], inp["rot_est"], inp["trans_est"],
                                                  inp["tgt_pose"], helpers.K_LM, means, stds, mode)
    np.testing.assert_allclose(pose, g[mode + "/pose"], rtol=0, atol=3e-7)
    np.testing.assert_allclose(rot, g[mode + "/rot"], rtol=0, atol=2e-6)
    np.testing.assert_allclose(trans, g[mode + "/trans"], rtol=1e-5, atol=2e-6)
    np.testing.assert_allclose(KT, g[mode + "/KT"], rtol=2e-6, atol=3e-4)


# ------------------------------------------------------------------ iteration-0 masks
def _mask_files(tmp_path):
    import cv2

    pairdb = helpers.write_pair_files(tmp_path, 41)
    labels = [cv2.imread(r["mask_gt_observed"], cv2.IMREAD_UNCHANGED) for r in pairdb]
    depths = [cv2.imread(r["depth_rendered"], cv2.IMREAD_UNCHANGED).astype(np.float32) / 1000
              for r in pairdb]
    return pairdb, labels, depths


def test_mask_preparation_restatement_vs_reference_loader(tmp_path):
    """oracle/prep_np.py (mask_dilate per pixel with the draws taken out, get_pair_mask for the
    configs' policies) against the reference's own get_pair_mask / mask_dilate
    (tests/golden/make_golden_masks.py): bit for bit, including the order of the random draws."""
    from oracle import prep_np

    g = np.load(os.path.join(GOLD, "masks_reference.npz"))
    pairdb, labels, depths = _mask_files(tmp_path)
    n = 2 * 480 * 640
    for phase, init in (("train", "box_gt"), ("test", "box_rendered")):
        for dil in (False, True):
            tag = "%s_%s_dil%d" % (phase, init, dil)
            np.random.seed(77)
            obs, gt, rend = [], [], []
            for rec, lab, d in zip(pairdb, labels, depths):
                draws = prep_np.draw_dilate(10) if dil else None
                if phase == "train":
                    o, m, r = prep_np.pair_masks_train_box_gt(lab, rec["mask_idx"], d, draws)
                else:
                    o, m, r = prep_np.pair_masks_test_box_rendered(d, draws)
                obs.append(o); gt.append(m); rend.append(r)
            ref_o = np.unpackbits(g[tag + "/mask_observed"])[:n].reshape(2, 480, 640)
            ref_g = np.unpackbits(g[tag + "/mask_gt_observed"])[:n].reshape(2, 480, 640)
            assert np.array_equal(np.stack(obs), ref_o), tag
            assert np.array_equal(np.stack(gt), ref_g), tag
            assert np.array_equal(np.stack(rend)[:, None, ::4, ::4], g[tag + "/mask_rendered"]), tag
    rng = np.random.default_rng(9)
    masks = (rng.random((12, 96, 128)) > 0.97).astype(np.float64)
    masks[3] *= 0.5
    masks[4, 10:20, 10:20] = 3.0
    np.random.seed(5)
    for m, ref in zip(masks, g["dilate/out"]):
        assert np.array_equal(prep_np.mask_dilate_from_draws(m, prep_np.draw_dilate(10)), ref)


def test_dilate_draws_of_the_product_follow_the_reference_rng_order():
    """prep.draw_dilate (host logic of the product) consumes np.random exactly like
    mask_dilate.py:19-47, i.e. like the restatement that is pinned above."""
    from mx_deepim_b200 import prep
    from oracle import prep_np

    np.random.seed(123)
    a = [prep.draw_dilate(10) for _ in range(200)]
    np.random.seed(123)
    b = [prep_np.draw_dilate(10) for _ in range(200)]
    assert a == b and any(0 in t for t in a) and max(max(t) for t in a) == 10
