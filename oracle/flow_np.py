"""oracle/flow_np.py -- TEST INFRASTRUCTURE ONLY (never on the product path).

* flow_kernel_cpu: ctypes front-end of oracle/flow_oracle.c, the CPU
  restatement of /root/reference/lib/flow_c/gpu_flow_kernel.cu:32-69.
* calc_flow: restatement of the numpy twin
  /root/reference/lib/pair_matching/flow.py:12-81 (+ backproject_camera,
  /root/reference/lib/utils/projection.py:47-67).  The reference function
  itself imports under numpy 2; tests/golden/make_golden.py runs it in the
  build container and stores its outputs (tests/golden/flow_calc_flow_*.npz),
  which pin this restatement bit-for-bit.
* reference_kernel_gpu: the reference .cu itself (oracle/_ref), GPU box only.
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

from . import build as _build
from . import pose_np

_lib = None


def _load():
    global _lib
    if _lib is None:
        path = _build.build_oracle()
        lib = ctypes.CDLL(path)
        lib.oracle_flow.restype = None
        lib.oracle_flow.argtypes = [ctypes.c_void_p] * 6 + [ctypes.c_int] * 4
        lib.oracle_max_threads.restype = ctypes.c_int
        _lib = lib
    return _lib


def max_threads():
    return int(_load().oracle_max_threads())


def flow_kernel_cpu(depth_src, depth_tgt, KT, Kinv, num_threads=0):
    """Same contract as gpu_flow (lib/flow_c/gpu_flow.pyx:24-41), computed on the CPU."""
    depth_src = np.ascontiguousarray(depth_src, dtype=np.float32)
    depth_tgt = np.ascontiguousarray(depth_tgt, dtype=np.float32)
    KT = np.ascontiguousarray(KT, dtype=np.float32)
    Kinv = np.ascontiguousarray(Kinv, dtype=np.float32)
    B, _, H, W = depth_src.shape
    flow = np.zeros((B, 2, H, W), np.float32)
    valid = np.zeros((B, 1, H, W), np.float32)
    if B:
        _load().oracle_flow(flow.ctypes.data, valid.ctypes.data, depth_src.ctypes.data,
                            depth_tgt.ctypes.data, KT.ctypes.data, Kinv.ctypes.data, B, H, W,
                            int(num_threads))
    return flow, valid


def backproject_camera(depth, intrinsic_matrix):
    """projection.py:47-67."""
    Kinv = np.linalg.inv(np.asarray(intrinsic_matrix))
    height, width = depth.shape[0], depth.shape[1]
    x, y = np.meshgrid(np.arange(width), np.arange(height))
    ones = np.ones((height, width), dtype=np.float32)
    x2d = np.stack((x, y, ones), axis=2).reshape(width * height, 3)
    R = np.dot(Kinv, x2d.transpose())
    X = np.multiply(np.tile(depth.reshape(1, width * height), (3, 1)), R)
    return np.array(X)


def calc_flow(depth_src, pose_src, pose_tgt, K, depth_tgt, thresh=3e-3, standard_rep=False):
    """lib/pair_matching/flow.py:12-81."""
    height, width = depth_src.shape[0], depth_src.shape[1]
    visible = np.zeros(depth_src.shape[:2]).flatten()
    X = backproject_camera(depth_src, intrinsic_matrix=K)
    transform = np.matmul(K, pose_np.se3_mul(pose_tgt, pose_np.se3_inverse(pose_src)))
    Xp = np.matmul(transform, np.append(X, np.ones([1, X.shape[1]], dtype=np.float32), axis=0))
    pz = Xp[2] + 1e-15
    pw = Xp[0] / pz
    ph = Xp[1] / pz
    valid_points = np.where(depth_src.flatten() != 0)[0]
    depth_proj_valid = pz[valid_points]
    pw_valid_raw = np.round(pw[valid_points]).astype(int)
    pw_valid = np.minimum(np.maximum(pw_valid_raw, 0), width - 1)
    ph_valid_raw = np.round(ph[valid_points]).astype(int)
    ph_valid = np.minimum(np.maximum(ph_valid_raw, 0), height - 1)
    p_within = np.logical_and(
        np.logical_and(pw_valid_raw >= 0, pw_valid_raw < width),
        np.logical_and(ph_valid_raw >= 0, ph_valid_raw < height))
    depth_tgt_valid = depth_tgt[ph_valid, pw_valid]
    p_within = np.logical_and(p_within, np.abs(depth_tgt_valid - depth_proj_valid) < thresh)
    p_valid = np.abs(depth_tgt_valid) > 1e-10
    fg_points = valid_points[np.logical_and(p_within, p_valid)]
    visible[fg_points] = 1
    visible = visible.reshape(depth_src.shape[:2])
    w_ori, h_ori = np.meshgrid(np.linspace(0, width - 1, width), np.linspace(0, height - 1, height))
    shape = depth_src.shape[:2]
    if standard_rep:
        flow = np.dstack([pw.reshape(shape) - w_ori, ph.reshape(shape) - h_ori])
    else:
        flow = np.dstack([ph.reshape(shape) - h_ori, pw.reshape(shape) - w_ori])
    flow[np.dstack([visible, visible]) != 1] = 0
    assert np.isnan(flow).sum() == 0
    X_valid = np.array([c[np.where(visible.flatten())] for c in X])
    return flow, visible, X_valid


# --------------------------------------------------------- reference on GPU
_ref = None


def reference_kernel_available():
    return os.path.exists(_build.REF_LIB)


def reference_kernel_gpu(depth_src, depth_tgt, KT, Kinv, device_id=0):
    """Run the reference's own `_flow` (gpu_flow_kernel.cu:82-148, compiled unmodified for
    sm_100a into oracle/_ref) -- host arrays in and out.  Needs a GPU."""
    global _ref
    if _ref is None:
        lib = ctypes.CDLL(_build.REF_LIB)
        fn = getattr(lib, "_Z5_flowPfS_S_S_S_S_iiii")  # C++ linkage: void _flow(float*,...,int,int,int,int)
        fn.restype = None
        fn.argtypes = [ctypes.c_void_p] * 6 + [ctypes.c_int] * 4
        _ref = fn
    depth_src = np.ascontiguousarray(depth_src, dtype=np.float32)
    depth_tgt = np.ascontiguousarray(depth_tgt, dtype=np.float32)
    KT = np.ascontiguousarray(KT, dtype=np.float32)
    Kinv = np.ascontiguousarray(Kinv, dtype=np.float32)
    B, _, H, W = depth_src.shape
    flow = np.zeros((B, 2, H, W), np.float32)
    valid = np.zeros((B, 1, H, W), np.float32)
    _ref(flow.ctypes.data, valid.ctypes.data, depth_src.ctypes.data, depth_tgt.ctypes.data,
         KT.ctypes.data, Kinv.ctypes.data, B, H, W, int(device_id))
    return flow, valid


def reference_kernel_dev(depth_src, depth_tgt, KT, Kinv, flow, valid, stream=0):
    """Launch the reference's unmodified `flow_kernel` (gpu_flow_kernel.cu:32-69, its own launch
    configuration :129-130) on DEVICE buffers -- arguments are raw device addresses (ints) of
    float32 arrays [B,1,H,W] x2, [B,3,4], [3,3], [B,2,H,W], [B,1,H,W] given as (ptr, shape) by the
    caller: kernel-only baseline G0.  No synchronisation."""
    lib = ctypes.CDLL(_build.REF_LIB)
    fn = lib.ref_flow_kernel_dev
    fn.restype = ctypes.c_int
    fn.argtypes = [ctypes.c_void_p] * 6 + [ctypes.c_int] * 3 + [ctypes.c_void_p]
    (ps, shape) = depth_src
    B, _, H, W = shape
    rc = fn(ps, depth_tgt, KT, Kinv, flow, valid, B, H, W, stream)
    if rc != 0:
        raise RuntimeError("ref_flow_kernel_dev: CUDA error %d" % rc)
