"""Zoom plugins on the GPU vs the numpy restatement (oracle/zoom_np.py).

Bars: crop boxes bit-exact; zoom factors bit-exact against the deterministic restatement;
resampled values within 1e-4 abs (north_star); rounded masks identical except where the
pre-round value sits within 1e-4 of the rounding threshold."""
import numpy as np
import pytest

import helpers
from oracle import zoom_np

pytestmark = pytest.mark.gpu
TOL = 1e-4
K_STR = str(helpers.K_LM.flatten())
PM_STR = str(helpers.PIXEL_MEANS.flatten())


def _t(arr, dev):
    import torch

    return torch.from_numpy(np.ascontiguousarray(arr)).to(dev)


def _custom(dev, **kw):
    import torch

    from mx_deepim_b200 import operator, operator_py  # noqa: F401

    kw = {k: (_t(v, dev) if isinstance(v, np.ndarray) and k not in ("K", "pixel_means") else v)
          for k, v in kw.items()}
    h = operator.Custom(**kw)
    torch.cuda.synchronize()
    return h


def _assert_rounded(got, ref, pre, thresh):
    """Rounded masks: identical wherever the pre-round value is not within TOL of `thresh`."""
    bad = got != ref
    assert not (bad & (np.abs(pre - thresh) > TOL)).any()
    assert bad.mean() < 1e-5


@pytest.mark.parametrize("batch,seed,empty", [(2, 5, ()), (4, 6, (1,))])
def test_zoom_mask(cuda_device, batch, seed, empty):
    d = helpers.make_zoom_inputs(batch, seed, empty_rendered=empty)
    r0, r1, r2, zf, bbox, pre = zoom_np.zoom_mask_forward(
        d["mask_observed"], d["mask_gt_observed"], d["mask_rendered"], d["src_pose"], helpers.K_LM,
        return_pre_round=True)
    h = _custom(cuda_device, op_type="ZoomMask", mask_observed=d["mask_observed"],
                mask_gt_observed=d["mask_gt_observed"], mask_rendered=d["mask_rendered"],
                src_pose=d["src_pose"], K=helpers.K_LM.flatten(), height=480, width=640)
    out = [o.cpu().numpy() for o in h.outputs]
    assert np.array_equal(out[3].view(np.uint32), zf.view(np.uint32))  # zoom factor bit-exact
    for got, ref, p in zip(out[:3], (r0, r1, r2), pre):
        _assert_rounded(got, ref, p, 0.5)
    # backward writes zeros
    import torch

    grads = h.backward([torch.ones_like(o) for o in h.outputs])
    assert all(float(g.abs().sum()) == 0.0 for g in grads)


def test_zoom_factor_bbox_bit_exact(cuda_device):
    import torch

    from mx_deepim_b200 import _lib
    from mx_deepim_b200.operator_py import _common as C

    d = helpers.make_zoom_inputs(5, 9, empty_rendered=(2,))
    _, _, _, zf, bbox = zoom_np.zoom_mask_forward(d["mask_observed"], d["mask_gt_observed"],
                                                  d["mask_rendered"], d["src_pose"], helpers.K_LM)
    zf_d = torch.empty((5, 4), dtype=torch.float32, device=cuda_device)
    bb_d = torch.empty((5, 8), dtype=torch.int32, device=cuda_device)
    C.zoom_factor(_t(d["mask_gt_observed"], cuda_device), _t(d["mask_rendered"], cuda_device),
                  _t(d["src_pose"], cuda_device), helpers.K_LM, None, zf_d, _lib.RC_ZF_MASK, bbox=bb_d)
    torch.cuda.synchronize()
    assert np.array_equal(bb_d.cpu().numpy().astype(np.int64), bbox)  # crop boxes bit-exact
    assert np.array_equal(zf_d.cpu().numpy().view(np.uint32), zf.view(np.uint32))


def test_zoom_image(cuda_device):
    d = helpers.make_zoom_inputs(2, 12)
    pm = helpers.PIXEL_MEANS[::-1]
    a, r, zf, bbox = zoom_np.zoom_image_forward(d["image_observed"], d["image_rendered"],
                                                d["src_pose"], helpers.K_LM, pm)
    h = _custom(cuda_device, op_type="ZoomImage", image_observed=d["image_observed"],
                image_rendered=d["image_rendered"], src_pose=d["src_pose"],
                K=helpers.K_LM.flatten(), height=480, width=640,
                pixel_means=helpers.PIXEL_MEANS.flatten())
    out = [o.cpu().numpy() for o in h.outputs]
    assert np.array_equal(out[2].view(np.uint32), zf.view(np.uint32))
    assert np.abs(out[0] - a).max() <= TOL and np.abs(out[1] - r).max() <= TOL


@pytest.mark.parametrize("spot", [False, True])
def test_zoom_image_with_factor(cuda_device, spot):
    d = helpers.make_zoom_inputs(3, 13)
    zf = zoom_np.zoom_mask_forward(d["mask_observed"], d["mask_gt_observed"], d["mask_rendered"],
                                   d["src_pose"], helpers.K_LM)[3]
    a, r = zoom_np.zoom_image_with_factor_forward(zf, d["image_observed"], d["image_rendered"],
                                                  helpers.PIXEL_MEANS[::-1], high_light_center=spot)
    h = _custom(cuda_device, op_type="ZoomImageWithFactor", zoom_factor=zf,
                image_observed=d["image_observed"], image_rendered=d["image_rendered"], height=480,
                width=640, pixel_means=helpers.PIXEL_MEANS.flatten(), high_light_center=spot)
    out = [o.cpu().numpy() for o in h.outputs]
    assert np.abs(out[0] - a).max() <= TOL
    assert np.abs(out[1] - r).max() <= TOL
    if spot:
        assert out[1][0, 0, 240, 320] == np.float32(255.0) - helpers.PIXEL_MEANS[2]


def test_zoom_depth_flow_mask_with_factor_trans(cuda_device):
    d = helpers.make_zoom_inputs(3, 14)
    zf = zoom_np.zoom_mask_forward(d["mask_observed"], d["mask_gt_observed"], d["mask_rendered"],
                                   d["src_pose"], helpers.K_LM)[3]
    # ZoomDepth
    a, r = zoom_np.zoom_depth_forward(zf, d["depth_observed"], d["depth_rendered"])
    h = _custom(cuda_device, op_type="ZoomDepth", zoom_factor=zf, depth_observed=d["depth_observed"],
                depth_rendered=d["depth_rendered"])
    assert np.abs(h.outputs[0].cpu().numpy() - a).max() <= TOL
    assert np.abs(h.outputs[1].cpu().numpy() - r).max() <= TOL
    # ZoomFlow forward
    zfl, zw, pre = zoom_np.zoom_flow_forward(zf, d["flow"], d["flow_weights"], return_pre_round=True)
    h = _custom(cuda_device, op_type="ZoomFlow", zoom_factor=zf, flow=d["flow"],
                flow_weights=d["flow_weights"], b_inv_zoom=False)
    got_f, got_w = [o.cpu().numpy() for o in h.outputs]
    assert np.abs(got_f - zfl).max() <= TOL * 4  # flow / wx: the 1e-4 px bar scaled by 1/wx <= 4
    _assert_rounded(got_w, zw, pre, 0.95)
    # ZoomFlow inverse of the zoomed flow
    inv = zoom_np.zoom_flow_forward(zf, zfl, b_inv_zoom=True)
    h = _custom(cuda_device, op_type="ZoomFlow", zoom_factor=zf, flow=zfl, b_inv_zoom=True)
    assert np.abs(h.outputs[0].cpu().numpy() - inv).max() <= TOL
    # ZoomMaskWithFactor, both directions
    for inv_flag in (False, True):
        ref, pre = zoom_np.zoom_mask_with_factor_forward(zf, d["mask_rendered"], inv_flag,
                                                         return_pre_round=True)
        h = _custom(cuda_device, op_type="ZoomMaskWithFactor", zoom_factor=zf,
                    mask=d["mask_rendered"], b_inv_zoom=inv_flag)
        _assert_rounded(h.outputs[0].cpu().numpy(), ref, pre, 0.5)
    # ZoomTrans forward / backward
    rng = np.random.default_rng(3)
    td = rng.normal(size=(3, 3)).astype(np.float32)
    for inv_flag in (False, True):
        for zg in (False, True):
            h = _custom(cuda_device, op_type="ZoomTrans", zoom_factor=zf, trans_delta=td,
                        b_inv_zoom=inv_flag, b_zoom_grad=zg)
            assert np.array_equal(h.outputs[0].cpu().numpy(),
                                  zoom_np.zoom_trans_forward(zf, td, inv_flag))
            import torch

            g = torch.from_numpy(td[::-1].copy()).to(cuda_device)
            gz, gt = h.backward([g])
            assert float(gz.abs().sum()) == 0.0
            assert np.array_equal(gt.cpu().numpy(),
                                  zoom_np.zoom_trans_backward(zf, td[::-1], inv_flag, zg))


def test_zoom_round_trip_property(cuda_device):
    """zoom -> inverse zoom of a flow field reproduces it inside the crop (recipe of
    zoom_flow.py:204-244 with zoom_factor (1.5, 1.5, 0.2, 0.2) replaced by a zoom-in)."""
    import torch

    d = helpers.make_zoom_inputs(2, 15)
    zf = np.tile(np.array([[0.5, 0.5, 0.1, -0.1]], np.float32), (2, 1))
    yy, xx = np.meshgrid(np.arange(480, dtype=np.float32), np.arange(640, dtype=np.float32),
                         indexing="ij")
    flow = np.tile(np.stack([0.01 * yy, 0.02 * xx])[None], (2, 1, 1, 1)).astype(np.float32)
    w = np.ones_like(flow)
    h = _custom(cuda_device, op_type="ZoomFlow", zoom_factor=zf, flow=flow, flow_weights=w,
                b_inv_zoom=False)
    h2 = _custom(cuda_device, op_type="ZoomFlow", zoom_factor=zf, flow=h.outputs[0].cpu().numpy(),
                 b_inv_zoom=True)
    back = h2.outputs[0].cpu().numpy()
    # crop window in source pixels: centre (tx+1)/2*W +- wx*W/2
    cx, cy = (0.1 + 1) / 2 * 640, (-0.1 + 1) / 2 * 480
    ys = slice(int(cy - 100), int(cy + 100))
    xs = slice(int(cx - 140), int(cx + 140))
    assert np.abs(back[:, :, ys, xs] - flow[:, :, ys, xs]).max() < 2e-2


def test_req_add_and_null(cuda_device):
    import torch

    d = helpers.make_zoom_inputs(1, 16)
    zf = np.array([[0.6, 0.6, 0.05, 0.0]], np.float32)
    ref = zoom_np.zoom_depth_forward(zf, d["depth_observed"], d["depth_rendered"])
    out = [torch.full((1, 1, 480, 640), 2.0, device=cuda_device) for _ in range(2)]
    _custom(cuda_device, op_type="ZoomDepth", zoom_factor=zf, depth_observed=d["depth_observed"],
            depth_rendered=d["depth_rendered"], out=out, req=["add", "null"])
    assert np.abs(out[0].cpu().numpy() - (ref[0] + 2.0)).max() <= TOL
    assert float((out[1] - 2.0).abs().max()) == 0.0


@pytest.mark.parametrize("height,width", [(152, 200), (250, 332)])
def test_zoom_other_frame_sizes(cuda_device, height, width):
    """The plugins take `height` / `width` kwargs (zoom_mask.py:143-150): a frame that is neither a
    multiple of the resampler's 128x8 tile nor 640 wide goes through the generic-width kernel;
    boxes and zoom factors stay bit-exact, samples within the 1e-4 bar."""
    rng = np.random.default_rng(height)
    B = 3
    K = np.array([[180.0, 0, width / 2 + 2.5], [0, 181.0, height / 2 - 1.5], [0, 0, 1]], np.float32)
    mask_gt = np.zeros((B, 1, height, width), np.float32)
    rend = np.zeros((B, 1, height, width), np.float32)
    src_pose = np.zeros((B, 3, 4), np.float32)
    for b in range(B):
        cy, cx = rng.integers(40, height - 40), rng.integers(50, width - 50)
        ry, rx = rng.integers(8, 30), rng.integers(8, 30)
        mask_gt[b, 0, cy - ry:cy + ry, cx - rx:cx + rx] = 1.0
        rend[b, 0, cy - ry + 3:cy + ry + 5, cx - rx - 4:cx + rx + 2] = rng.uniform(0.5, 1.0)
        src_pose[b, :, :3] = np.eye(3)
        z = 0.8
        src_pose[b, :, 3] = [(cx - K[0, 2]) / K[0, 0] * z, (cy - K[1, 2]) / K[1, 1] * z, z]
    mask_obs = mask_gt.copy()
    mask_obs[:, :, ::7] = 0
    r0, r1, r2, zf, bbox, pre = zoom_np.zoom_mask_forward(mask_obs, mask_gt, rend, src_pose, K,
                                                          height=height, width=width,
                                                          return_pre_round=True)
    h = _custom(cuda_device, op_type="ZoomMask", mask_observed=mask_obs, mask_gt_observed=mask_gt,
                mask_rendered=rend, src_pose=src_pose, K=K.flatten(), height=height, width=width)
    out = [o.cpu().numpy() for o in h.outputs]
    assert np.array_equal(out[3].view(np.uint32), zf.view(np.uint32))
    for got, ref, p in zip(out[:3], (r0, r1, r2), pre):
        bad = got != ref
        assert not (bad & (np.abs(p - 0.5) > TOL)).any()
    pm = helpers.PIXEL_MEANS
    img_o = rng.uniform(-120, 130, size=(B, 3, height, width)).astype(np.float32)
    img_r = rng.uniform(-120, 130, size=(B, 3, height, width)).astype(np.float32)
    ref_o, ref_r = zoom_np.zoom_image_with_factor_forward(zf, img_o, img_r, pm[::-1])
    h = _custom(cuda_device, op_type="ZoomImageWithFactor", zoom_factor=zf, image_observed=img_o,
                image_rendered=img_r, height=height, width=width, pixel_means=pm.flatten())
    out = [o.cpu().numpy() for o in h.outputs]
    assert np.abs(out[0] - ref_o).max() <= TOL and np.abs(out[1] - ref_r).max() <= TOL


# ---------------------------------------------------------------- independent sampler pin
PIN_CASES = [(0.32, 0.32, 0.05, -0.1), (0.11, 0.11, -0.4, 0.3), (1.7, 1.7, 0.3, -0.2),
             (0.4, 0.4, 0.85, -0.9), (0.999, 0.999, 0.0, 0.0), (0.25, 0.25, 0.2, 0.1)]


@pytest.mark.parametrize("inverse", [False, True])
def test_zoom_sample_kernel_vs_torch_grid_sample(cuda_device, inverse):
    """rc_zoom_sample_dev against an INDEPENDENT implementation of the same operator pair:
    torch.nn.functional.affine_grid + grid_sample (align_corners=True, zeros padding) on the
    GPU, in float64 (the semantics pinned in tests/test_oracle_cpu.py) -- a checker only, never a
    product dependency.  Bars = the float32 coordinate-rounding bounds measured for the float32
    restatement (tests/test_oracle_cpu.py, x2 margin): the kernel must be as close to the exact
    operator as MXNet's own float32 arithmetic is.  Smooth field and image-range noise; zoom-in,
    zoom-out, off-image windows, inverse zoom (zoom_flow.py:35-44).  And against the float32
    restatement itself at the north_star bar (1e-4 abs)."""
    import torch
    import torch.nn.functional as F

    from mx_deepim_b200.operator_py import _common as C

    B = len(PIN_CASES)
    rng = np.random.default_rng(321)
    smooth = np.fromfunction(lambda b, c, y, x: np.sin(x / 37.0) + np.cos(y / 23.0) + c + 0 * b,
                             (B, 2, 480, 640)).astype(np.float32)
    noise = rng.uniform(-255, 255, size=(B, 2, 480, 640)).astype(np.float32)
    zf = np.array(PIN_CASES, np.float32)
    zf_d = _t(zf, cuda_device)
    if inverse:
        aff = np.array([[float(v) for v in zoom_np.inverse_affine(z, 480, 640)] for z in zf], np.float64)
    else:
        aff = zf.astype(np.float64)
    theta = torch.tensor([[[a[0], 0.0, a[2]], [0.0, a[1], a[3]]] for a in aff], dtype=torch.float64,
                         device=cuda_device)
    for name, data, bar in (("smooth", smooth, 4.2e-4), ("noise", noise, 0.13)):
        src = _t(data, cuda_device)
        dst = torch.empty_like(src)
        C.zoom_sample(zf_d, [C.Group(src, dst)], 480, 640, inverse=inverse)
        torch.cuda.synchronize()
        grid = F.affine_grid(theta, [B, 2, 480, 640], align_corners=True)
        ref = F.grid_sample(src.double(), grid, mode="bilinear", padding_mode="zeros",
                            align_corners=True)
        err = (dst.double() - ref).abs().amax(dim=(1, 2, 3)).cpu().numpy()
        assert err.max() <= bar, (name, inverse, err)
        if name == "smooth" and not inverse:
            assert err[[0, 1, 5]].max() <= 4e-6, err  # zoom-in inside the image: the production case
        # and the float32 restatement (MXNet's arithmetic) at the north_star bar
        mine = zoom_np.sample_batch(data, zf, inverse=inverse)
        d = np.abs(dst.cpu().numpy() - mine)
        assert d.max() <= (1e-4 if name == "noise" else 2e-6), (name, inverse, d.max())
