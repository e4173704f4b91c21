"""Shared launch helpers for the zoom / Transform3D plugins (ctypes -> C ABI)."""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from .. import _lib


def _req_flag(req):
    """Map an MXNet write request to (skip, accumulate)."""
    if req == "null":
        return True, 0
    if req in ("write", "inplace"):
        return False, 0
    if req == "add":
        return False, 1
    raise ValueError("unknown req %r" % (req,))


def check_tensor(name, t, ndim=None, dtype=torch.float32, shape=None):
    """A tensor crosses into the C ABI as a bare pointer: it has to be a CUDA tensor of the
    declared dtype and C-contiguous (a transposed or sliced view would be read with the wrong
    strides, silently)."""
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise TypeError("%s must be a CUDA torch.Tensor (no CPU fallback)" % name)
    if t.dtype != dtype:
        raise TypeError("%s must be %s, got %s" % (name, dtype, t.dtype))
    if not t.is_contiguous():
        raise ValueError("%s must be contiguous" % name)
    if ndim is not None and t.dim() != ndim:
        raise ValueError("%s must have %d dims, got %d" % (name, ndim, t.dim()))
    if shape is not None and tuple(t.shape) != tuple(shape):
        raise ValueError("%s must have shape %s, got %s" % (name, tuple(shape), tuple(t.shape)))


class Group(object):
    """One rc_zoom_group: a [B,C,H,W] source resampled into a [B,C,H,W] destination."""

    def __init__(self, src, dst, pre_op=_lib.RC_PRE_NONE, post_op=_lib.RC_POST_NONE, req="write",
                 mean=None, broadcast_src=False):
        self.src, self.dst, self.pre_op, self.post_op, self.req, self.mean = (
            src, dst, pre_op, post_op, req, mean)
        self.broadcast_src = broadcast_src


def zoom_sample(zoom_factor, groups, height, width, inverse=False, stream=None):
    """One launch of the fused bilinear resampler over every group in `groups`."""
    check_tensor("zoom_factor", zoom_factor, 2)
    B = zoom_factor.shape[0]
    arr = (_lib.rc_zoom_group * _lib.RC_ZOOM_MAX_GROUPS)()
    n = 0
    for g in groups:
        skip, acc = _req_flag(g.req)
        if skip:
            continue
        check_tensor("zoom source", g.src, 4)
        check_tensor("zoom destination", g.dst, 4)
        src_shape = tuple(g.src.shape)
        if g.broadcast_src and src_shape[1] == 1:
            src_shape = (src_shape[0], g.dst.shape[1]) + src_shape[2:]
        if src_shape != tuple(g.dst.shape) or g.src.shape[0] != B or \
                g.src.shape[2] != height or g.src.shape[3] != width:
            raise ValueError("zoom group shape mismatch: src %s dst %s expected [%d,C,%d,%d]"
                             % (tuple(g.src.shape), tuple(g.dst.shape), B, height, width))
        if n >= _lib.RC_ZOOM_MAX_GROUPS:
            raise ValueError("too many zoom groups in one launch")
        e = arr[n]
        e.src = g.src.data_ptr()
        e.dst = g.dst.data_ptr()
        e.channels = g.dst.shape[1]
        e.pre_op = g.pre_op
        e.post_op = g.post_op
        e.flags = (_lib.RC_ZOOM_ACCUMULATE if acc else 0) | (
            _lib.RC_ZOOM_SRC_BROADCAST if g.broadcast_src else 0)
        mean = [0.0] * 4 if g.mean is None else [float(v) for v in g.mean] + [0.0] * (4 - len(g.mean))
        for i in range(4):
            e.mean[i] = mean[i]
        n += 1
    if n == 0:
        return
    lib = _lib.load()
    with torch.cuda.device(zoom_factor.device):
        st = lib.rc_zoom_sample_dev(zoom_factor.data_ptr(), 1 if inverse else 0, arr, n, B,
                                    height, width, _lib.stream_ptr(stream))
    _lib.check(st, "rc_zoom_sample_dev")


_CONST_CACHE = {}


def device_const(values, device):
    """Small float32 constant (K, pixel means) resident on `device`, cached by value."""
    arr = np.asarray(values, dtype=np.float32).reshape(-1)
    key = (arr.tobytes(), str(device))
    t = _CONST_CACHE.get(key)
    if t is None:
        t = torch.from_numpy(arr.copy()).to(device)
        _CONST_CACHE[key] = t
    return t


_WS_CACHE = {}


def _workspace(batch, device, stream_handle):
    """Zeroed bbox workspace.  The kernel leaves it zeroed, so the buffer is reused -- but only
    by launches that are ordered behind each other: one buffer per (device, B, stream).  Two
    CustomOps with the same batch on different engine streams therefore never share one
    (they would race on the atomics and on the ticket counter)."""
    key = (str(device), int(batch), int(stream_handle))
    ws = _WS_CACHE.get(key)
    if ws is None:
        ws = torch.zeros((batch, _lib.RC_ZF_WS_INTS), dtype=torch.int32, device=device)
        _WS_CACHE[key] = ws
    return ws


def zoom_factor(obs, rend, src_pose, K, pixel_means, zoom_factor_out, mode, bbox=None,
                stream=None):
    """Crop box + zoom factor on device (rc_zoom_factor_dev)."""
    check_tensor("obs", obs, 4)
    check_tensor("rend", rend, 4)
    check_tensor("src_pose", src_pose, 3)
    check_tensor("zoom_factor", zoom_factor_out, 2)
    B, C, H, W = obs.shape
    dev = obs.device
    Kd = device_const(K, dev)
    pm = device_const(pixel_means, dev) if pixel_means is not None else None
    lib = _lib.load()
    with torch.cuda.device(dev):
        sp = _lib.stream_ptr(stream)
        ws = _workspace(B, dev, sp)
        st = lib.rc_zoom_factor_dev(obs.data_ptr(), rend.data_ptr(), src_pose.data_ptr(),
                                    Kd.data_ptr(), pm.data_ptr() if pm is not None else None,
                                    zoom_factor_out.data_ptr(),
                                    bbox.data_ptr() if bbox is not None else None, ws.data_ptr(),
                                    B, C, H, W, mode, sp)
        if st != _lib.RC_OK:
            # a failed / aborted launch may leave partial boxes or a ticket behind: drop the
            # buffer so that the next call starts from a freshly zeroed one
            _WS_CACHE.pop((str(dev), int(B), int(sp)), None)
    _lib.check(st, "rc_zoom_factor_dev")


def fill_zero(t, req):
    """Every zoom backward writes zeros (e.g. zoom_mask.py:136-140)."""
    skip, acc = _req_flag(req)
    if skip or acc:  # "add" of zero is a no-op
        return
    lib = _lib.load()
    with torch.cuda.device(t.device):
        st = lib.rc_fill_zero_dev(t.data_ptr(), t.numel(), _lib.stream_ptr())
    _lib.check(st, "rc_fill_zero_dev")
