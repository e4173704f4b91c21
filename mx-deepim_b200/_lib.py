"""ctypes binding of librendercompare.so (C ABI declared in include/rendercompare.h).

This is the only place the shared library is opened.  There is no CPU
fallback: if the library is missing or a call fails, an exception is raised.
The binding plays the role of the reference's Cython stub
(/root/reference/lib/flow_c/gpu_flow.pyx:12-16).
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_float, c_int, c_int32, c_int64, c_void_p

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG_DIR, "librendercompare.so")

RC_OK = 0
# enums mirrored from include/rendercompare.h
RC_ZF_MASK, RC_ZF_IMAGE = 0, 1
RC_PRE_NONE, RC_PRE_THRESH02, RC_PRE_ADD_MEAN = 0, 1, 2
(RC_POST_NONE, RC_POST_ROUND, RC_POST_ROUND_M045, RC_POST_DIV_WX, RC_POST_MUL_WX,
 RC_POST_SUB_MEAN, RC_POST_SPOT_SUB_MEAN) = range(7)
RC_ROT_MODEL, RC_ROT_CAMERA, RC_ROT_CAMERA_NEW, RC_ROT_NAIVE = range(4)
RC_PM_L1, RC_PM_L2, RC_PM_SMOOTH_L1 = range(3)
RC_ZOOM_MAX_GROUPS = 8
RC_ZOOM_MAX_PLANES = 16
RC_ZF_WS_INTS = 10

ROT_COORD = {"model": RC_ROT_MODEL, "camera": RC_ROT_CAMERA, "camera_new": RC_ROT_CAMERA_NEW,
             "naive": RC_ROT_NAIVE}
PM_LOSS_TYPE = {"l1": RC_PM_L1, "l2": RC_PM_L2, "smooth_l1": RC_PM_SMOOTH_L1}


class RenderCompareError(RuntimeError):
    """A librendercompare entry point returned a non-zero status."""


class rc_zoom_group(ctypes.Structure):
    _fields_ = [
        ("src", c_void_p),
        ("dst", c_void_p),
        ("channels", c_int32),
        ("pre_op", c_int32),
        ("post_op", c_int32),
        ("accumulate", c_int32),
        ("mean", c_float * 4),
    ]


# name -> (restype, argtypes); every symbol include/rendercompare.h declares.
_P = c_void_p
SIGNATURES = {
    "rc_version": (c_int, []),
    "rc_status_string": (c_char_p, [c_int]),
    "rc_last_cuda_error": (c_char_p, []),
    "rc_last_status": (c_int, []),
    "rc_device_sm_count": (c_int, [c_int, POINTER(c_int)]),
    "_flow": (None, [_P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int]),
    "rc_flow_host": (c_int, [_P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int]),
    "rc_flow_dev": (c_int, [_P, _P, _P, _P, _P, _P, c_int, c_int, c_int, _P]),
    "rc_zoom_factor_dev": (c_int, [_P, _P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int,
                                   c_int, _P]),
    "rc_zoom_sample_dev": (c_int, [_P, c_int, POINTER(rc_zoom_group), c_int, c_int, c_int, c_int,
                                   _P]),
    "rc_zoom_trans_dev": (c_int, [_P, _P, _P, c_int, c_int, c_int, c_int, _P]),
    "rc_fill_zero_dev": (c_int, [_P, c_int64, _P]),
    "rc_transform3d_fwd_dev": (c_int, [_P, _P, _P, _P, POINTER(c_float), POINTER(c_float), c_int,
                                       _P, c_int, c_int, c_int, _P]),
    "rc_transform3d_bwd_dev": (c_int, [_P, _P, _P, _P, _P, POINTER(c_float), POINTER(c_float),
                                       c_int, _P, _P, c_int, c_int, c_int, _P]),
    "rc_pm_loss_fwd_dev": (c_int, [_P, _P, _P, _P, _P, _P, POINTER(c_float), POINTER(c_float),
                                   c_int, c_int, c_float, c_float, _P, _P, _P, c_int, c_int, _P]),
    "rc_pm_loss_bwd_dev": (c_int, [_P, _P, _P, _P, _P, _P, POINTER(c_float), POINTER(c_float),
                                   c_int, c_int, c_float, c_float, c_float, _P, _P, c_int, c_int,
                                   _P]),
}

_lib = None


def load() -> ctypes.CDLL:
    """Open the shared library once and attach prototypes.  Raises if absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RenderCompareError(
            "librendercompare.so is not built (%s). Build it with "
            "`python -c 'import __graft_entry__ as g; g.build()'` from the repo root; "
            "there is no CPU fallback." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here = header/library mismatch
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(status: int, what: str) -> None:
    if status != RC_OK:
        lib = load()
        msg = lib.rc_status_string(status).decode()
        if status == 2:
            msg += ": " + lib.rc_last_cuda_error().decode()
        raise RenderCompareError("%s failed: %s" % (what, msg))


def f3(values) -> ctypes.Array:
    """Three host floats (T_means / T_stds kwargs)."""
    arr = (c_float * 3)()
    vals = [0.0, 0.0, 0.0] if values is None else [float(v) for v in values]
    for i in range(3):
        arr[i] = vals[i]
    return arr


def stream_ptr(stream=None) -> int:
    """cudaStream_t of `stream` (torch.cuda.Stream) or of torch's current stream."""
    import torch

    s = torch.cuda.current_stream() if stream is None else stream
    return int(s.cuda_stream)
