"""Ground-truth optical flow -- drop-in for the reference's ``lib/flow_c``.

Mirrors
  /root/reference/lib/flow_c/flow.py:17-21      gpu_flow_wrapper(device_id)
  /root/reference/lib/flow_c/gpu_flow.pyx:24-41 gpu_flow(depth_src, depth_tgt, KT, Kinv, device_id)

``gpu_flow`` keeps the reference contract: float32 numpy arrays in
(``[B,1,H,W]``, ``[B,1,H,W]``, ``[B,3,4]``, ``[3,3]``), numpy arrays out
(``flow [B,2,H,W]`` with channel 0 = delta-row, channel 1 = delta-col, and
``valid [B,1,H,W]``).  The Cython stub rejected wrong dtypes / ranks with
ValueError ("Buffer dtype mismatch" / "Buffer has wrong number of
dimensions"); the same exception type is raised here.

``gpu_flow_dev`` is the fast path: torch CUDA tensors in and out, no host
round trip, asynchronous on the current stream.
"""
from __future__ import annotations

import numpy as np

from . import _lib


def _check_buffer(name, arr, ndim):
    if not isinstance(arr, np.ndarray):
        raise TypeError("Argument '%s' has incorrect type (expected numpy.ndarray, got %s)"
                        % (name, type(arr).__name__))
    if arr.dtype != np.float32:
        raise ValueError("Buffer dtype mismatch, expected 'float32_t' but got '%s' (%s)"
                         % (arr.dtype, name))
    if arr.ndim != ndim:
        raise ValueError("Buffer has wrong number of dimensions (expected %d, got %d) (%s)"
                         % (ndim, arr.ndim, name))


def gpu_flow(depth_src, depth_tgt, KT, Kinv, device_id=0):
    """flow, valid = gpu_flow(depth_src, depth_tgt, KT, Kinv, device_id)  (host arrays)."""
    _check_buffer("depth_src", depth_src, 4)
    _check_buffer("depth_tgt", depth_tgt, 4)
    _check_buffer("KT", KT, 3)
    _check_buffer("Kinv", Kinv, 2)
    batch_size, _, height, width = depth_src.shape
    flow = np.zeros([batch_size, 2, height, width], dtype=np.float32)
    valid = np.zeros([batch_size, 1, height, width], dtype=np.float32)
    if batch_size > 0:  # gpu_flow.pyx:36
        lib = _lib.load()
        # The Cython stub passes &arr[0,0,0,0]; it therefore assumes C-contiguity.
        depth_src = np.ascontiguousarray(depth_src)
        depth_tgt = np.ascontiguousarray(depth_tgt)
        KT = np.ascontiguousarray(KT)
        Kinv = np.ascontiguousarray(Kinv)
        st = lib.rc_flow_host(flow.ctypes.data, valid.ctypes.data, depth_src.ctypes.data,
                              depth_tgt.ctypes.data, KT.ctypes.data, Kinv.ctypes.data,
                              batch_size, height, width, int(device_id))
        _lib.check(st, "rc_flow_host")
    return flow, valid


def gpu_flow_wrapper(device_id):
    """Reference lib/flow_c/flow.py:17-21."""

    def _flow(depth_src, depth_tgt, KT, Kinv):
        return gpu_flow(depth_src, depth_tgt, KT, Kinv, device_id)

    return _flow


def gpu_flow_dev(depth_src, depth_tgt, KT, Kinv, flow=None, valid=None, stream=None):
    """Device fast path: torch.cuda float32 tensors, same shapes as gpu_flow."""
    import torch

    for name, t in (("depth_src", depth_src), ("depth_tgt", depth_tgt), ("KT", KT), ("Kinv", Kinv)):
        if not (isinstance(t, torch.Tensor) and t.is_cuda and t.dtype == torch.float32
                and t.is_contiguous()):
            raise ValueError("%s must be a contiguous float32 CUDA tensor" % name)
    B, _, H, W = depth_src.shape
    if flow is None:
        flow = torch.empty((B, 2, H, W), dtype=torch.float32, device=depth_src.device)
    if valid is None:
        valid = torch.empty((B, 1, H, W), dtype=torch.float32, device=depth_src.device)
    lib = _lib.load()
    with torch.cuda.device(depth_src.device):
        st = lib.rc_flow_dev(depth_src.data_ptr(), depth_tgt.data_ptr(), KT.data_ptr(),
                             Kinv.data_ptr(), flow.data_ptr(), valid.data_ptr(), B, H, W,
                             _lib.stream_ptr(stream))
    _lib.check(st, "rc_flow_dev")
    return flow, valid
