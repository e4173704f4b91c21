"""oracle/zoom_np.py -- TEST INFRASTRUCTURE ONLY (never on the product path).

numpy restatement of the reference zoom plugins
  /root/reference/deepim/operator_py/zoom_mask.py:29-134
  /root/reference/deepim/operator_py/zoom_image.py:26-119
  /root/reference/deepim/operator_py/zoom_image_with_factor.py:31-75
  /root/reference/deepim/operator_py/zoom_depth.py:24-50
  /root/reference/deepim/operator_py/zoom_flow.py:28-77
  /root/reference/deepim/operator_py/zoom_mask_with_factor.py:29-68
  /root/reference/deepim/operator_py/zoom_trans.py:22-76

Two parts with different provenance:

* the crop-box / zoom-factor block is LIFTED from the reference's numpy code
  (zoom_mask.py:55-105); only `asnumpy()` inputs are replaced by arrays.  Where
  the reference mixes an np.float32 scalar with a Python / int64 scalar, its
  numpy (1.x, "legacy" value-based promotion) evaluates in float64; numpy >= 2
  (NEP 50) would stay float32.  The float64 casts below make the numpy-1.x
  behaviour explicit so the result does not depend on the installed numpy.

* GridGenerator(affine) + BilinearSampler live in Apache MXNet (pip
  `mxnet-cu90`, "tested under mxnet 1.2.0", requirements.txt:5, README.md:81-98)
  which is neither vendored nor installable here.  They are restated from the
  published MXNet 1.2 operators (src/operator/grid_generator-inl.h,
  src/operator/bilinear_sampler.cu): normalised grid -1 + i*2/(n-1), affine
  by a [2x3] GEMM, x_real = (g+1)*(n-1)/2, zero contribution from taps outside
  [0, n-1], float32 tap weights and a float64-promoted blend.

PINNED for everything the plugins compute in Python / numpy: the reference's
own operator files run unmodified under a numpy `mxnet` stand-in
(tests/golden/make_golden_ops.py, mx_stub.py) and every function below
reproduces their outputs bit-exactly (tests/test_golden_ops_cpu.py).
PARITY UNPINNED for the two MXNet built-ins (GridGenerator, BilinearSampler):
the reference ships no golden vectors and no asserting test for them
(SURVEY.md section 4); the restatement is cross-checked against
torch.nn.functional.grid_sample(align_corners=True, padding_mode="zeros") in
tests/test_oracle_cpu.py.
"""
from __future__ import annotations

import numpy as np

import contextlib

f32 = np.float32
f64 = np.float64

# Scalar promotion of the reference's mixed np.float32 / Python-scalar expressions:
#   "legacy" -- numpy 1.x, which the reference was written for: evaluated in float64 (default;
#               what the product implements);
#   "nep50"  -- numpy >= 2: stays float32.  Used only to compare with the golden vectors that
#               tests/golden/make_golden_ops.py produces by running the reference's own plugin
#               files under this container's numpy 2.x.
PROMOTION = "legacy"


@contextlib.contextmanager
def promotion(mode):
    global PROMOTION
    assert mode in ("legacy", "nep50")
    old, PROMOTION = PROMOTION, mode
    try:
        yield
    finally:
        PROMOTION = old


# ----------------------------------------------------------------- rounding
def mx_round(x):
    """mx.nd.round: C roundf, half away from zero (float32 in, float32 out)."""
    x = np.asarray(x, dtype=f32)
    a = np.floor(np.abs(x).astype(f64) + 0.5)
    return (np.sign(x).astype(f64) * a).astype(f32)


def fma32(a, b, c):
    """float32 fused multiply-add, exact: product and sum are exact in float64
    for the operand ranges used here (|a| < 2^11 integer, 24-bit b, c = -1)."""
    return (np.asarray(a, f64) * np.asarray(b, f64) + np.asarray(c, f64)).astype(f32)


# -------------------------------------------------------------- zoom factor
def _bbox(valid):
    """zoom_mask.py:55-62: min/max index of the non-empty columns / rows."""
    x_max = np.max(valid, axis=0)
    y_max = np.max(valid, axis=1)
    nz_x = np.nonzero(x_max)[0]
    nz_y = np.nonzero(y_max)[0]
    return nz_x, nz_y


def zoom_factor_from_valid(valid_real, valid_rendered, K, src_pose, height, width):
    """One sample of zoom_mask.py:51-119 / zoom_image.py:41-118.

    valid_real, valid_rendered: bool [H,W]; K float32 [3,3]; src_pose float32 [3,4].
    Returns (zoom_factor float32 [4], bbox int [8] = real x0,x1,y0,y1, rendered x0,x1,y0,y1
    with (x0 > x1) marking an empty rendered set)."""
    K = np.asarray(K, dtype=f32)
    src_pose = np.asarray(src_pose, dtype=f32)
    nz_x, nz_y = _bbox(valid_real)
    obj_real_start_x = np.min(nz_x)  # raises on an empty observed mask, like the reference
    obj_real_end_x = np.max(nz_x)
    obj_real_start_y = np.min(nz_y)
    obj_real_end_y = np.max(nz_y)
    obj_real_c_x = (obj_real_start_x + obj_real_end_x) * 0.5
    obj_real_c_y = (obj_real_start_y + obj_real_end_y) * 0.5

    nz_x, nz_y = _bbox(valid_rendered)
    # np.dot(self.K, src_pose[:, 3]) on float32 operands (:72); evaluated here as
    # sequential float32 multiply-adds so the value does not depend on the BLAS build
    t = src_pose[:, 3]
    obj_rendered_c = np.array(
        [f32(f32(f32(K[r, 0] * t[0]) + f32(K[r, 1] * t[1])) + f32(K[r, 2] * t[2])) for r in range(3)],
        dtype=f32)
    obj_rendered_c_x = f32(obj_rendered_c[0] / obj_rendered_c[2])
    obj_rendered_c_y = f32(obj_rendered_c[1] / obj_rendered_c[2])
    if len(nz_x) == 0 or len(nz_y) == 0:
        obj_rendered_start_x = obj_real_start_x
        obj_rendered_end_x = obj_real_end_x
        obj_rendered_start_y = obj_real_start_y
        obj_rendered_end_y = obj_real_end_y
        zoom_c_x = f64(obj_real_c_x)
        zoom_c_y = f64(obj_real_c_y)
        rbox = (2 ** 31 - 1, -1, 2 ** 31 - 1, -1)
    else:
        obj_rendered_start_x = np.min(nz_x)
        obj_rendered_end_x = np.max(nz_x)
        obj_rendered_start_y = np.min(nz_y)
        obj_rendered_end_y = np.max(nz_y)
        zoom_c_x = f64(obj_rendered_c_x)  # np.float32 op int64 -> float64 from here on
        zoom_c_y = f64(obj_rendered_c_y)
        rbox = (obj_rendered_start_x, obj_rendered_end_x, obj_rendered_start_y, obj_rendered_end_y)

    left_dist = max(zoom_c_x - obj_rendered_start_x, zoom_c_x - obj_real_start_x)
    right_dist = max(obj_rendered_end_x - zoom_c_x, obj_real_end_x - zoom_c_x)
    up_dist = max(zoom_c_y - obj_rendered_start_y, zoom_c_y - obj_real_start_y)
    down_dist = max(obj_real_end_y - zoom_c_y, obj_rendered_end_y - zoom_c_y)
    crop_height = np.max([0.75 * right_dist, 0.75 * left_dist, up_dist, down_dist]) * 1.4 * 2

    wx = crop_height / height
    wy = crop_height / height
    tx = zoom_c_x / width * 2 - 1
    ty = zoom_c_y / height * 2 - 1
    if PROMOTION == "nep50" and rbox[1] >= 0:
        # NEP 50: np.float32 centre with Python ints stays float32 (three float32 roundings)
        tx = f32(f32(f32(obj_rendered_c_x / f32(width)) * f32(2)) - f32(1))
        ty = f32(f32(f32(obj_rendered_c_y / f32(height)) * f32(2)) - f32(1))
    zf = np.array([wx, wy, tx, ty], dtype=f64).astype(f32)  # zoom_factor_array[b, i] = ... (:116-119)
    bbox = np.array([obj_real_start_x, obj_real_end_x, obj_real_start_y, obj_real_end_y, *rbox],
                    dtype=np.int64)
    return zf, bbox


def valid_from_masks(mask_gt_observed, mask_rendered):
    """zoom_mask.py:36-47.  Inputs [B,C,H,W] float32.  Returns bool [B,H,W] x2 and the
    thresholded rendered mask (float32 0/1) that is then resampled."""
    valid_real = np.sum(np.asarray(mask_gt_observed, f32), axis=1) > 0.3
    m = np.array(mask_rendered, dtype=f32, copy=True)
    hi = m > 0.2
    m[hi] = 1
    m[~hi & (m <= 0.2)] = 0
    valid_rendered = np.sum(m, axis=1) > 0.3
    return valid_real, valid_rendered, m


def valid_from_images(image_observed, image_rendered, pixel_means):
    """zoom_image.py:31-37.  pixel_means float32 [1,3,1,1] (already reversed)."""
    pm = np.asarray(pixel_means, f32).reshape(1, -1, 1, 1)
    a = (np.asarray(image_observed, f32) + pm).astype(f32)
    b = (np.asarray(image_rendered, f32) + pm).astype(f32)

    def seq_sum(v):  # np.sum over a short strided axis adds sequentially in float32
        s = v[:, 0].copy()
        for c in range(1, v.shape[1]):
            s = (s + v[:, c]).astype(f32)
        return s

    return seq_sum(a) > 0.01, seq_sum(b) > 0.01


# ------------------------------------------------------- grid + sampler
def grid_generator_affine(wx, wy, tx, ty, height, width):
    """MXNet GridGenerator(transform_type="affine") for the matrix [[wx,0,tx],[0,wy,ty]].

    Returns (gx, gy) float32 [H,W] in normalised [-1,1] coordinates."""
    wx, wy, tx, ty = f32(wx), f32(wy), f32(tx), f32(ty)
    xn = fma32(np.arange(width, dtype=f32), f32(2.0 / (width - 1)), -1.0)
    yn = fma32(np.arange(height, dtype=f32), f32(2.0 / (height - 1)), -1.0)
    gx = ((wx * xn).astype(f32) + tx).astype(f32)
    gy = ((wy * yn).astype(f32) + ty).astype(f32)
    return np.broadcast_to(gx[None, :], (height, width)), np.broadcast_to(gy[:, None], (height, width))


def bilinear_sampler(data, gx, gy):
    """MXNet BilinearSampler for one sample.  data [C,H,W] float32; gx, gy [Ho,Wo]."""
    data = np.asarray(data, dtype=f32)
    C, H, W = data.shape
    x_real = (((gx + f32(1)).astype(f32) * f32(W - 1)).astype(f32) / f32(2)).astype(f32)
    y_real = (((gy + f32(1)).astype(f32) * f32(H - 1)).astype(f32) / f32(2)).astype(f32)
    x0f = np.floor(x_real)
    y0f = np.floor(y_real)
    x0 = x0f.astype(np.int64)
    y0 = y0f.astype(np.int64)
    wyt = (f32(1) - (y_real - y0f).astype(f32)).astype(f32)
    wxl = (f32(1) - (x_real - x0f).astype(f32)).astype(f32)

    def tap(yy, xx):
        ok = (xx >= 0) & (xx <= W - 1) & (yy >= 0) & (yy <= H - 1)
        v = data[:, np.clip(yy, 0, H - 1), np.clip(xx, 0, W - 1)]
        return np.where(ok[None], v, f32(0)).astype(f32)

    tl, tr = tap(y0, x0), tap(y0, x0 + 1)
    bl, br = tap(y0 + 1, x0), tap(y0 + 1, x0 + 1)
    one_m_x = 1.0 - wxl.astype(f64)
    one_m_y = 1.0 - wyt.astype(f64)
    t1 = ((tl * wyt).astype(f32) * wxl).astype(f32).astype(f64)
    t2 = (tr * wyt).astype(f32).astype(f64) * one_m_x
    t3 = bl.astype(f64) * one_m_y * wxl.astype(f64)
    t4 = br.astype(f64) * one_m_y * one_m_x
    return (((t1 + t2) + t3) + t4).astype(f32)


def sample_affine_f64(data, wx, wy, tx, ty):
    """The SAME formulas as grid_generator_affine + bilinear_sampler, every operation in float64:
    the semantic twin of the float32 restatement (which has to follow MXNet's float32 rounding).
    tests/test_oracle_cpu.py pins it to torch.nn.functional.grid_sample(align_corners=True,
    padding_mode="zeros") in float64 to 1e-9, i.e. the SEMANTICS (normalised grid, pixel mapping,
    floor / weight / zero-tap rules) are pinned independently of the float32 arithmetic."""
    data = np.asarray(data, dtype=f64)
    C, H, W = data.shape
    xn = np.arange(W, dtype=f64) * (2.0 / (W - 1)) - 1.0
    yn = np.arange(H, dtype=f64) * (2.0 / (H - 1)) - 1.0
    gx = np.broadcast_to((f64(wx) * xn + f64(tx))[None, :], (H, W))
    gy = np.broadcast_to((f64(wy) * yn + f64(ty))[:, None], (H, W))
    x_real = (gx + 1.0) * (W - 1) / 2.0
    y_real = (gy + 1.0) * (H - 1) / 2.0
    x0f, y0f = np.floor(x_real), np.floor(y_real)
    x0, y0 = x0f.astype(np.int64), y0f.astype(np.int64)
    wyt = 1.0 - (y_real - y0f)
    wxl = 1.0 - (x_real - x0f)

    def tap(yy, xx):
        ok = (xx >= 0) & (xx <= W - 1) & (yy >= 0) & (yy <= H - 1)
        v = data[:, np.clip(yy, 0, H - 1), np.clip(xx, 0, W - 1)]
        return np.where(ok[None], v, 0.0)

    return (tap(y0, x0) * wyt * wxl + tap(y0, x0 + 1) * wyt * (1.0 - wxl) +
            tap(y0 + 1, x0) * (1.0 - wyt) * wxl + tap(y0 + 1, x0 + 1) * (1.0 - wyt) * (1.0 - wxl))


def inverse_affine(zf, height, width):
    """zoom_flow.py:35-44 / zoom_mask_with_factor.py:43-52 (float64 scalars, numpy 1.x)."""
    if PROMOTION == "nep50":
        return _inverse_affine_f32(zf, height, width)
    wx_in, wy_in, tx_in, ty_in = [f64(v) for v in np.asarray(zf, f32)]
    wx = 1 / wx_in
    wy = 1 / wy_in
    crop_width = wx_in * width
    crop_height = wy_in * height
    obj_real_c_x = tx_in * 0.5 * width + 0.5 * width
    obj_real_c_y = ty_in * 0.5 * height + 0.5 * height
    tx = (width * 0.5 - obj_real_c_x) / crop_width * 2
    ty = (height * 0.5 - obj_real_c_y) / crop_height * 2
    return f32(wx), f32(wy), f32(tx), f32(ty)


def _inverse_affine_f32(zf, height, width):
    """The same expressions with every operation rounded to float32 (NEP 50 promotion)."""
    wx_in, wy_in, tx_in, ty_in = [f32(v) for v in np.asarray(zf, f32)]
    w, h = f32(width), f32(height)
    wx = f32(f32(1) / wx_in)
    wy = f32(f32(1) / wy_in)
    crop_width = f32(wx_in * w)
    crop_height = f32(wy_in * h)
    obj_real_c_x = f32(f32(f32(tx_in * f32(0.5)) * w) + f32(0.5 * width))
    obj_real_c_y = f32(f32(f32(ty_in * f32(0.5)) * h) + f32(0.5 * height))
    tx = f32(f32(f32(f32(width * 0.5) - obj_real_c_x) / crop_width) * f32(2))
    ty = f32(f32(f32(f32(height * 0.5) - obj_real_c_y) / crop_height) * f32(2))
    return wx, wy, tx, ty


def sample_batch(data, zoom_factor, inverse=False):
    """Resample [B,C,H,W] with per-sample zoom factors [B,4]."""
    data = np.asarray(data, dtype=f32)
    B, C, H, W = data.shape
    out = np.empty_like(data)
    for b in range(B):
        if inverse:
            wx, wy, tx, ty = inverse_affine(zoom_factor[b], H, W)
        else:
            wx, wy, tx, ty = [f32(v) for v in zoom_factor[b]]
        gx, gy = grid_generator_affine(wx, wy, tx, ty, H, W)
        out[b] = bilinear_sampler(data[b], gx, gy)
    return out


# ----------------------------------------------------------- the operators
def zoom_mask_forward(mask_observed, mask_gt_observed, mask_rendered, src_pose, K, height=480,
                      width=640, return_pre_round=False):
    """ZoomMaskOperator.forward (zoom_mask.py:29-134)."""
    valid_real, valid_rendered, rendered01 = valid_from_masks(mask_gt_observed, mask_rendered)
    B = valid_real.shape[0]
    zf = np.zeros((B, 4), dtype=f32)
    bbox = np.zeros((B, 8), dtype=np.int64)
    for b in range(B):
        zf[b], bbox[b] = zoom_factor_from_valid(valid_real[b], valid_rendered[b], K, src_pose[b],
                                                height, width)
    pre = [sample_batch(m, zf) for m in (mask_observed, mask_gt_observed, rendered01)]
    outs = [mx_round(p) for p in pre]
    if return_pre_round:
        return outs[0], outs[1], outs[2], zf, bbox, pre
    return outs[0], outs[1], outs[2], zf, bbox


def zoom_image_forward(image_observed, image_rendered, src_pose, K, pixel_means, height=480,
                       width=640):
    """ZoomImageOperator.forward (zoom_image.py:26-119); pixel_means as the Prop stores it."""
    pm = np.asarray(pixel_means, f32).reshape(1, -1, 1, 1)
    valid_real, valid_rendered = valid_from_images(image_observed, image_rendered, pm)
    B = valid_real.shape[0]
    zf = np.zeros((B, 4), dtype=f32)
    bbox = np.zeros((B, 8), dtype=np.int64)
    for b in range(B):
        zf[b], bbox[b] = zoom_factor_from_valid(valid_real[b], valid_rendered[b], K, src_pose[b],
                                                height, width)
    a = (sample_batch((np.asarray(image_observed, f32) + pm).astype(f32), zf) - pm).astype(f32)
    r = (sample_batch((np.asarray(image_rendered, f32) + pm).astype(f32), zf) - pm).astype(f32)
    return a, r, zf, bbox


def center_map(batch_size, height, width, spot_radius=5):
    """zoom_image_with_factor.py:36-45."""
    cm = np.zeros([batch_size, 3, height, width], dtype=f32)
    sx = int(np.floor(float(width) / 2 - spot_radius))
    ex = int(np.ceil(float(width) / 2 + spot_radius))
    sy = int(np.floor(float(height) / 2 - spot_radius))
    ey = int(np.ceil(float(height) / 2 + spot_radius))
    cm[:, 0, sy:ey, sx:ex] = 255.0
    return cm


def zoom_image_with_factor_forward(zoom_factor, image_observed, image_rendered, pixel_means,
                                   high_light_center=False):
    """ZoomImageWithFactorOperator.forward (zoom_image_with_factor.py:31-75).
    Returns (zoom_image_observed, zoom_image_rendered)."""
    pm = np.asarray(pixel_means, f32).reshape(1, -1, 1, 1)
    B, _, H, W = image_observed.shape
    real = sample_batch((np.asarray(image_observed, f32) + pm).astype(f32), zoom_factor)
    rend = sample_batch((np.asarray(image_rendered, f32) + pm).astype(f32), zoom_factor)
    if high_light_center:
        rend = np.maximum(rend, center_map(B, H, W))
    return (real - pm).astype(f32), (rend - pm).astype(f32)


def zoom_depth_forward(zoom_factor, depth_observed, depth_rendered):
    """ZoomDepthOperator.forward (zoom_depth.py:24-50)."""
    return sample_batch(depth_observed, zoom_factor), sample_batch(depth_rendered, zoom_factor)


def zoom_flow_forward(zoom_factor, flow, flow_weights=None, b_inv_zoom=False,
                      return_pre_round=False):
    """ZoomFlowOperator.forward (zoom_flow.py:28-77)."""
    zf = np.asarray(zoom_factor, f32)
    out = sample_batch(flow, zf, inverse=b_inv_zoom)
    for b in range(out.shape[0]):
        wx, wy = zf[b, 0], zf[b, 1]
        assert wx == wy, "wx and wy should be equal"
        out[b] = (out[b] * wx).astype(f32) if b_inv_zoom else (out[b] / wx).astype(f32)
    if b_inv_zoom:
        return out
    pre = sample_batch(flow_weights, zf)
    w = mx_round((pre - f32(0.45)).astype(f32))
    if return_pre_round:
        return out, w, pre
    return out, w


def zoom_mask_with_factor_forward(zoom_factor, mask, b_inv_zoom=False, return_pre_round=False):
    """ZoomMaskWithFactorOperator.forward (zoom_mask_with_factor.py:29-68)."""
    m = np.array(mask, dtype=f32, copy=True)
    hi = m > 0.2
    m[hi] = 1
    m[~hi & (m <= 0.2)] = 0
    pre = sample_batch(m, zoom_factor, inverse=b_inv_zoom)
    if return_pre_round:
        return mx_round(pre), pre
    return mx_round(pre)


def zoom_trans_forward(zoom_factor, trans_delta, b_inv_zoom=False):
    """ZoomTransOperator.forward (zoom_trans.py:22-48): np.float32 scalar arithmetic."""
    zf = np.asarray(zoom_factor, f32)
    td = np.asarray(trans_delta, f32)
    out = np.zeros(td.shape)
    for b in range(td.shape[0]):
        wx = zf[b][0]
        wy = zf[b][0]
        dx, dy, dz = td[b]
        if b_inv_zoom:
            out[b, 0], out[b, 1] = f32(dx * wx), f32(dy * wy)
        else:
            out[b, 0], out[b, 1] = f32(dx / wx), f32(dy / wy)
        out[b, 2] = dz
    return out.astype(f32)


def zoom_trans_backward(zoom_factor, out_grad, b_inv_zoom=False, b_zoom_grad=False):
    """ZoomTransOperator.backward (zoom_trans.py:50-76)."""
    zf = np.asarray(zoom_factor, f32)
    g = np.asarray(out_grad, f32)
    out = np.zeros(g.shape)
    for b in range(g.shape[0]):
        wx = zf[b][0]
        gx, gy, gz = g[b]
        if b_zoom_grad:
            if b_inv_zoom:
                gx, gy = f32(gx * wx), f32(gy * wx)
            else:
                gx, gy = f32(gx / wx), f32(gy / wx)
        out[b] = (gx, gy, gz)
    return out.astype(f32)
