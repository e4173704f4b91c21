"""Iteration-0 input preparation on the device.

Mirrors the loader-side functions that feed the render-and-compare path, with the same names
and argument meaning, over CUDA tensors:

    transform(im, pixel_means)                    /root/reference/lib/utils/image.py:709-720
    decode_depth(raw, depth_factor, mask, idx)    image.py:200-206, :229-260
    calc_flow(depth_src, pose_src, pose_tgt, K, depth_tgt, thresh, standard_rep)
                                                  /root/reference/lib/pair_matching/flow.py:12-81
    get_pair_mask(label, idx, depth_rendered, phase, init_mask, dilate)   image.py:272-492
    mask_dilate(mask, draws) / draw_dilate()      /root/reference/lib/utils/mask_dilate.py:10-58
    get_point_cloud_model(points_obj, num_sample) image.py:559-589
    get_point_cloud_observed(points, pose)        image.py:592-596

`calc_flow` keeps the reference's numpy interface (numpy in, numpy out, float64 results) and
has a batched device twin `calc_flow_batch`.  There is no CPU fallback: every function needs
the CUDA library and a CUDA device.
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from . import _lib
from . import pose as P


def _dev(device):
    if not torch.cuda.is_available():
        raise _lib.RenderCompareError("input preparation needs a CUDA device (no CPU fallback)")
    return torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)


def _check(name, t, dtype, shape=None):
    """Every tensor that crosses into the C ABI is passed as a bare pointer: it has to be a
    CUDA tensor of the declared dtype, C-contiguous (a transposed or sliced view would be read
    with the wrong strides -- silently), and of the declared shape."""
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise TypeError("%s must be a CUDA torch.Tensor (no CPU fallback)" % name)
    if t.dtype != dtype:
        raise TypeError("%s must be %s, got %s" % (name, dtype, t.dtype))
    if not t.is_contiguous():
        raise ValueError("%s must be contiguous" % name)
    if shape is not None and tuple(t.shape) != tuple(shape):
        raise ValueError("%s must have shape %s, got %s" % (name, tuple(shape), tuple(t.shape)))


def transform_batch(bgr_u8, pixel_means, out=None, stream=None):
    """bgr_u8: CUDA uint8 [B,H,W,3] (BGR as cv2.imread delivers).  pixel_means: 3 values in
    B,G,R order (config.network.PIXEL_MEANS).  Returns float32 [B,3,H,W] RGB - means."""
    if bgr_u8.dtype != torch.uint8 or bgr_u8.dim() != 4 or bgr_u8.shape[3] != 3:
        raise TypeError("bgr_u8 must be a uint8 [B,H,W,3] tensor")
    if not bgr_u8.is_cuda or not bgr_u8.is_contiguous():
        raise ValueError("bgr_u8 must be a contiguous CUDA tensor")
    B, H, W, _ = bgr_u8.shape
    if out is None:
        out = torch.empty((B, 3, H, W), dtype=torch.float32, device=bgr_u8.device)
    _check("out", out, torch.float32, (B, 3, H, W))
    means = (ctypes.c_double * 3)(*[float(v) for v in np.asarray(pixel_means, np.float64).reshape(-1)[:3]])
    with torch.cuda.device(bgr_u8.device):
        st = _lib.load().rc_prep_image_dev(bgr_u8.data_ptr(), out.data_ptr(), means, B, H, W,
                                           _lib.stream_ptr(stream))
    _lib.check(st, "rc_prep_image_dev")
    return out


def transform(im, pixel_means, device=None):
    """Drop-in for image.transform: im [H,W,3] BGR uint8 (numpy) -> numpy [1,3,H,W]."""
    dev = _dev(device)
    t = torch.from_numpy(np.ascontiguousarray(im, np.uint8)[None]).to(dev)
    out = transform_batch(t, pixel_means)
    torch.cuda.synchronize(dev)
    return out.cpu().numpy()


def decode_depth(raw_u16, depth_factor, mask=None, mask_idx=None, out=None, stream=None):
    """raw_u16: CUDA [B,H,W] uint16 (torch.uint16) or int16 view of the PNG depth.
    mask: CUDA uint8 [B,H,W] + mask_idx int32 [B] to zero the depth outside the instance
    (image.py:244).  Returns float32 [B,1,H,W] metres."""
    if raw_u16.element_size() != 2 or raw_u16.dim() != 3:
        raise TypeError("raw_u16 must be a 16-bit [B,H,W] tensor")
    if not raw_u16.is_cuda or not raw_u16.is_contiguous():
        raise ValueError("raw_u16 must be a contiguous CUDA tensor")
    B, H, W = raw_u16.shape
    if out is None:
        out = torch.empty((B, 1, H, W), dtype=torch.float32, device=raw_u16.device)
    _check("out", out, torch.float32, (B, 1, H, W))
    if (mask is None) != (mask_idx is None):
        raise ValueError("mask and mask_idx go together")
    if mask is not None:
        _check("mask", mask, torch.uint8, (B, H, W))
        _check("mask_idx", mask_idx, torch.int32, (B,))
    with torch.cuda.device(raw_u16.device):
        st = _lib.load().rc_prep_depth_dev(
            raw_u16.data_ptr(), mask.data_ptr() if mask is not None else None,
            mask_idx.data_ptr() if mask_idx is not None else None, out.data_ptr(),
            float(depth_factor), B, H, W, _lib.stream_ptr(stream))
    _lib.check(st, "rc_prep_depth_dev")
    return out


def decode_mask(mask_u8, mask_idx, out=None, stream=None):
    """get_pair_mask's instance selection (image.py:300-304): CUDA uint8 [B,H,W] + int32 [B]
    -> float32 [B,1,H,W] with 1.0 where mask == mask_idx."""
    if mask_u8.dtype != torch.uint8 or mask_u8.dim() != 3:
        raise TypeError("mask_u8 must be a uint8 [B,H,W] tensor")
    B, H, W = mask_u8.shape
    _check("mask_u8", mask_u8, torch.uint8)
    _check("mask_idx", mask_idx, torch.int32, (B,))
    if out is None:
        out = torch.empty((B, 1, H, W), dtype=torch.float32, device=mask_u8.device)
    _check("out", out, torch.float32, (B, 1, H, W))
    with torch.cuda.device(mask_u8.device):
        st = _lib.load().rc_prep_mask_dev(mask_u8.data_ptr(), mask_idx.data_ptr(), out.data_ptr(),
                                          B, H, W, _lib.stream_ptr(stream))
    _lib.check(st, "rc_prep_mask_dev")
    return out


def mask_from_depth(depth, thresh=0.2, keep_small=False, out=None, stream=None):
    """depth CUDA float32 (metres, any shape, numel % 4 == 0) -> depth > thresh ? 1 :
    (keep_small ? depth : 0).  keep_small=True is get_pair_mask's mask_rendered
    (image.py:487-492), False the binary mask the init policies box from (:335-337, :446-447)."""
    _check("depth", depth, torch.float32)
    if out is None:
        out = torch.empty_like(depth)
    _check("out", out, torch.float32, depth.shape)
    with torch.cuda.device(depth.device):
        st = _lib.load().rc_mask_from_depth_dev(depth.data_ptr(), out.data_ptr(), float(thresh),
                                                1 if keep_small else 0, depth.numel(),
                                                _lib.stream_ptr(stream))
    _lib.check(st, "rc_mask_from_depth_dev")
    return out


def draw_dilate(max_thickness=10, rng=np.random):
    """The random draws of ONE mask_dilate call in the reference's order (mask_dilate.py:19-47):
    `direction`, then a thickness for every block that direction does not skip.  Returns the
    four thicknesses (rows down, rows up, columns right, columns left), 0 = block skipped."""
    direction = rng.randint(10)
    t = [0, 0, 0, 0]
    for i, skip in enumerate(([0, 1, 4], [1, 2, 5], [2, 3, 6], [0, 3, 7])):
        if direction not in skip:
            t[i] = rng.randint(max_thickness) + 1
    return t


def mask_dilate(mask, thickness, out=None, stream=None):
    """mask_dilate (lib/utils/mask_dilate.py:10-58) on the device for a batch: mask CUDA float32
    [B,1,H,W]; thickness [B,4] (draw_dilate per pair; tensor or array-like).  Returns [B,1,H,W]."""
    _check("mask", mask, torch.float32)
    if mask.dim() != 4 or mask.shape[1] != 1:
        raise ValueError("mask must be [B,1,H,W]")
    B, _, H, W = mask.shape
    if not isinstance(thickness, torch.Tensor):
        thickness = torch.from_numpy(np.ascontiguousarray(thickness, np.int32).reshape(B, 4)).to(mask.device)
    _check("thickness", thickness, torch.int32, (B, 4))
    if out is None:
        out = torch.empty_like(mask)
    _check("out", out, torch.float32, mask.shape)
    if out.data_ptr() == mask.data_ptr():
        raise ValueError("out must not alias mask")
    with torch.cuda.device(mask.device):
        st = _lib.load().rc_mask_dilate_dev(mask.data_ptr(), thickness.data_ptr(), out.data_ptr(),
                                            B, H, W, _lib.stream_ptr(stream))
    _lib.check(st, "rc_mask_dilate_dev")
    return out


def get_pair_mask(label_u8, mask_idx, depth_rendered, phase="train", init_mask=None,
                  dilate=False, max_thickness=10, rng=np.random, stream=None):
    """get_pair_mask (lib/utils/image.py:272-492) for a batch on the device, for the policies
    the shipped configs use: phase "train" with TRAIN.INIT_MASK "box_gt", phase "test" with
    TEST.INIT_MASK "box_rendered"; `dilate` = *.MASK_DILATE (draws come from `rng` pair by pair,
    in the reference's order).

    label_u8 CUDA uint8 [B,H,W] + mask_idx int32 [B] (train only, may be None for test);
    depth_rendered CUDA float32 [B,1,H,W] in metres (prep.decode_depth).
    Returns (mask_observed, mask_gt_observed, mask_rendered), each float32 [B,1,H,W]; at test
    time mask_gt_observed IS mask_observed (image.py:480-481)."""
    from .pipeline import mask_box

    if init_mask is None:
        init_mask = "box_gt" if phase == "train" else "box_rendered"
    if phase == "train":
        if init_mask != "box_gt":
            raise Exception("Unknown mask type: {}".format(init_mask))  # image.py:343
        gt = decode_mask(label_u8, mask_idx, stream=stream)
        obs, _ = mask_box(gt, stream=stream)
    else:
        if init_mask != "box_rendered":
            raise Exception("Unknown init mask type: {}".format(init_mask))  # image.py:467-470
        binary = mask_from_depth(depth_rendered, 0.2, keep_small=False, stream=stream)
        obs, _ = mask_box(binary, stream=stream)
        gt = None
    if dilate:
        B = depth_rendered.shape[0]
        draws = [draw_dilate(max_thickness, rng) for _ in range(B)]
        obs = mask_dilate(obs, draws, stream=stream)
    rend = mask_from_depth(depth_rendered, 0.2, keep_small=True, stream=stream)
    return obs, (obs if gt is None else gt), rend


def get_point_cloud_model(points_obj, num_sample, rng=np.random):
    """image.py:559-589 without the file cache: shuffle, keep <= num_sample points, zero-pad.
    Returns ([1,3,num_sample] points, [1,3,num_sample] weights) float64 numpy, like the
    reference.  Host side: one-off per batch, driven by numpy's global RNG."""
    points_obj = np.asarray(points_obj)
    num_points = points_obj.shape[0]
    num_keep = min(num_points, num_sample)
    keep_idx = np.arange(num_points)
    rng.shuffle(keep_idx)
    keep_idx = keep_idx[:num_keep]
    points_sample = np.zeros((num_sample, 3))
    points_sample[:num_keep, :] = points_obj[keep_idx, :]
    points_weights = np.zeros((num_sample, 3))
    points_weights[:num_keep, :] = 1
    return np.expand_dims(points_sample.T, axis=0), np.expand_dims(points_weights.T, axis=0)


def get_point_cloud_observed(points_model, pose_observed, out=None, stream=None):
    """image.py:592-596 batched on the device: points_model CUDA float32 [B,3,N], pose_observed
    CUDA float32 [B,3,4] -> [B,3,N]."""
    _check("points_model", points_model, torch.float32)
    if points_model.dim() != 3 or points_model.shape[1] != 3:
        raise ValueError("points_model must be [B,3,N]")
    B, _, N = points_model.shape
    _check("pose_observed", pose_observed, torch.float32, (B, 3, 4))
    if out is None:
        out = torch.empty_like(points_model)
    _check("out", out, torch.float32, (B, 3, N))
    with torch.cuda.device(points_model.device):
        st = _lib.load().rc_points_transform_dev(points_model.data_ptr(), pose_observed.data_ptr(),
                                                 out.data_ptr(), B, N, _lib.stream_ptr(stream))
    _lib.check(st, "rc_points_transform_dev")
    return out


def make_transform(K, pose_src, pose_tgt):
    """flow.py:31: np.matmul(K, se3_mul(pose_tgt, se3_inverse(pose_src))) -- float64 K times
    the float32-rounded relative transform (projection.py:12-43)."""
    return np.matmul(np.asarray(K), P.se3_mul(pose_tgt, P.se3_inverse(pose_src)))


def calc_flow_batch(depth_src, depth_tgt, transform, K, thresh=3e-3, standard_rep=False,
                    want_points=False, capacity=None, stream=None):
    """Batched calc_flow on the device.  depth_src / depth_tgt: CUDA float32 [B,1,H,W];
    transform: CUDA float64 [B,3,4] (make_transform per pair); K: 3x3 (host).
    Returns (flow [B,2,H,W] f32, visible [B,1,H,W] f32, x_valid [B,3,capacity] f32 or None,
    counts [B] int32 or None).  No synchronisation."""
    _check("depth_src", depth_src, torch.float32)
    if depth_src.dim() != 4 or depth_src.shape[1] != 1:
        raise ValueError("depth_src must be [B,1,H,W]")
    B, _, H, W = depth_src.shape
    dev = depth_src.device
    _check("depth_tgt", depth_tgt, torch.float32, (B, 1, H, W))
    if transform.dtype != torch.float64 or tuple(transform.shape) != (B, 3, 4):
        raise TypeError("transform must be float64 [B,3,4]")
    _check("transform", transform, torch.float64, (B, 3, 4))
    lib = _lib.load()
    flow = torch.empty((B, 2, H, W), dtype=torch.float32, device=dev)
    visible = torch.empty((B, 1, H, W), dtype=torch.float32, device=dev)
    x_valid = counts = ws = None
    cap = 0
    if want_points:
        cap = H * W if capacity is None else int(capacity)
        x_valid = torch.empty((B, 3, cap), dtype=torch.float32, device=dev)
        counts = torch.empty((B,), dtype=torch.int32, device=dev)
        ws = torch.empty((B * lib.rc_calc_flow_blocks(H, W),), dtype=torch.int32, device=dev)
    Kinv = np.ascontiguousarray(np.linalg.inv(np.asarray(K)), np.float64).reshape(-1)
    kin = (ctypes.c_double * 9)(*[float(v) for v in Kinv])
    with torch.cuda.device(dev):
        st = lib.rc_calc_flow_dev(
            depth_src.data_ptr(), depth_tgt.data_ptr(), transform.data_ptr(), kin, float(thresh),
            1 if standard_rep else 0, flow.data_ptr(), visible.data_ptr(),
            x_valid.data_ptr() if x_valid is not None else None,
            counts.data_ptr() if counts is not None else None,
            ws.data_ptr() if ws is not None else None, cap, B, H, W, _lib.stream_ptr(stream))
    _lib.check(st, "rc_calc_flow_dev")
    return flow, visible, x_valid, counts


def calc_flow(depth_src, pose_src, pose_tgt, K, depth_tgt, thresh=3e-3, standard_rep=False,
              device=None):
    """Drop-in for lib.pair_matching.flow.calc_flow (same arguments and return values:
    flow [H,W,2], visible [H,W], X_valid [3,Nv]) computed on the device."""
    dev = _dev(device)
    H, W = depth_src.shape[:2]
    ds = torch.from_numpy(np.ascontiguousarray(depth_src, np.float32).reshape(1, 1, H, W)).to(dev)
    dt = torch.from_numpy(np.ascontiguousarray(depth_tgt, np.float32).reshape(1, 1, H, W)).to(dev)
    T = torch.from_numpy(np.ascontiguousarray(make_transform(K, pose_src, pose_tgt), np.float64)
                         .reshape(1, 3, 4)).to(dev)
    flow, visible, xv, counts = calc_flow_batch(ds, dt, T, K, thresh, standard_rep, want_points=True)
    torch.cuda.synchronize(dev)
    n = int(counts[0].item())
    return (flow[0].permute(1, 2, 0).double().cpu().numpy(), visible[0, 0].double().cpu().numpy(),
            xv[0, :, :n].double().cpu().numpy())
