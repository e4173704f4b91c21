"""oracle/raster_np.py -- TEST INFRASTRUCTURE ONLY (never on the product path).

numpy/ctypes front-end of oracle/raster_oracle.c: the CPU restatement of the
reference's OpenGL renderers.

  my_compute_calib_proj, _get_view_mtx : LIFTED from
      /root/reference/lib/render_glumpy/render_py_multi.py:152-178 (pure numpy).
  render_py            : Render_Py.render(cls_idx, r, t, r_type, K)            (:112-147)
  render_py_light_modelnet : Render_Py_Light_ModelNet_Multi.render(...)
      /root/reference/lib/render_glumpy/render_py_light_modelnet_multi.py:149-231

PARITY UNPINNED (see raster_oracle.c header): no GL on this box, no rendered goldens in the
reference.  Cross-oracle pin used instead (north_star): the rendered depth must be consistent
with the flow oracle -- see tests/test_oracle_cpu.py::test_raster_oracle_cross_checks; the projection / view matrices are
pinned against the reference renderer classes themselves (tests/golden/make_golden_render.py).
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import build as _build
from . import pose_np

_lib = None


def _load():
    global _lib
    if _lib is None:
        lib = ctypes.CDLL(_build.build_oracle())
        lib.oracle_render.restype = None
        P = ctypes.c_void_p
        lib.oracle_render.argtypes = [P, P, P, P, ctypes.c_int, ctypes.c_int, P, ctypes.c_int,
                                      ctypes.c_int, P, P, ctypes.c_int, ctypes.c_int,
                                      ctypes.c_double, ctypes.c_double, ctypes.c_int, P, P,
                                      ctypes.c_double, P, P, P]
        _lib = lib
    return _lib


def my_compute_calib_proj(K, w, h, zNear, zFar):
    """render_py_multi.py:152-169."""
    u0 = K[0, 2] + 0.5
    v0 = K[1, 2] + 0.5
    fu = K[0, 0]
    fv = K[1, 1]
    L = +(u0) * zNear / -fu
    T = +(v0) * zNear / fv
    R = -(w - u0) * zNear / -fu
    B = -(h - v0) * zNear / fv
    proj = np.zeros((4, 4))
    proj[0, 0] = 2 * zNear / (R - L)
    proj[1, 1] = 2 * zNear / (T - B)
    proj[2, 2] = -(zFar + zNear) / (zFar - zNear)
    proj[2, 0] = (R + L) / (R - L)
    proj[2, 1] = (T + B) / (T - B)
    proj[2, 3] = -1.0
    proj[3, 2] = -(2 * zFar * zNear) / (zFar - zNear)
    return proj


def _get_view_mtx(R, t):
    """render_py_multi.py:171-178."""
    u_view = np.eye(4, dtype=np.float32)
    u_view[:3, :3], u_view[:3, 3] = R, np.asarray(t).squeeze()
    yz_flip = np.eye(4, dtype=np.float32)
    yz_flip[1, 1], yz_flip[2, 2] = -1, -1
    u_view = yz_flip.dot(u_view)
    u_view = u_view.T
    return u_view


def _render(mesh, R, t, K, width, height, zNear, zFar, mode, light_pos=None, light_int=None,
            brightness_ratio=0.7, want_tri=False):
    verts = np.ascontiguousarray(mesh["vertices"], dtype=np.float64)
    uv = np.ascontiguousarray(mesh["uv"], dtype=np.float64) if mesh.get("uv") is not None else None
    nrm = (np.ascontiguousarray(mesh["normals"], dtype=np.float64)
           if mesh.get("normals") is not None else None)
    tris = np.ascontiguousarray(mesh["triangles"], dtype=np.int32)
    tex = np.ascontiguousarray(mesh["texture"], dtype=np.uint8)
    # K is float32 in the reference (config.py:143-156); its numpy (1.x) evaluates
    # `K[0, 2] + 0.5` etc. in float64 (np.float32 scalar op Python float), numpy >= 2 would stay
    # in float32.  Passing the same values as float64 reproduces the numpy-1.x arithmetic.
    proj = np.ascontiguousarray(
        my_compute_calib_proj(np.asarray(K, dtype=np.float64), width, height, zNear, zFar),
        dtype=np.float64)
    # the uniform is float32 on the GL side; keep its float32 values, evaluate in float64
    view = np.ascontiguousarray(_get_view_mtx(R, t), dtype=np.float64)
    bgr = np.zeros((height, width, 3), np.float32)
    depth = np.zeros((height, width), np.float32)
    tri = np.zeros((height, width), np.int32) if want_tri else None
    lp = np.ascontiguousarray(light_pos if light_pos is not None else [0, 0, 0], dtype=np.float64)
    li = np.ascontiguousarray(light_int if light_int is not None else [1, 1, 1], dtype=np.float64)
    _load().oracle_render(
        verts.ctypes.data, uv.ctypes.data if uv is not None else None,
        nrm.ctypes.data if nrm is not None else None, tris.ctypes.data, verts.shape[0],
        tris.shape[0], tex.ctypes.data, tex.shape[0], tex.shape[1], proj.ctypes.data,
        view.ctypes.data, width, height, float(zNear), float(zFar), mode, lp.ctypes.data,
        li.ctypes.data, float(brightness_ratio), bgr.ctypes.data, depth.ctypes.data,
        tri.ctypes.data if tri is not None else None)
    return (bgr, depth, tri) if want_tri else (bgr, depth)


def render_py(mesh, r, t, K, r_type="quat", width=640, height=480, zNear=0.25, zFar=6.0,
              want_tri=False):
    """Render_Py.render: returns (bgr [H,W,3] float32 0..255, depth [H,W] float32 metres)."""
    R = pose_np.quat2mat(r) if r_type == "quat" else np.asarray(r)
    return _render(mesh, R, t, K, width, height, zNear, zFar, 0, want_tri=want_tri)


def render_py_light_modelnet(mesh, r, t, light_position, light_intensity, K, r_type="quat",
                             width=640, height=480, zNear=0.25, zFar=6.0, brightness_ratio=0.7):
    """Render_Py_Light_ModelNet_Multi.render: returns (bgr uint8 [H,W,3], depth float32)."""
    R = pose_np.quat2mat(r) if r_type == "quat" else np.asarray(r)
    bgr, depth = _render(mesh, R, t, K, width, height, zNear, zFar, 1, light_position,
                         light_intensity, brightness_ratio)
    return bgr.astype(np.uint8), depth


def render_py_light(mesh, r, t, light_position, light_intensity, K, r_type="quat", width=640,
                    height=480, zNear=0.25, zFar=6.0, brightness_ratio=0.4):
    """Render_Py_Light.render / Render_Py_Light_MultiProgram.render
    (/root/reference/lib/render_glumpy/render_py_light.py:136-213,
    render_py_light_multi_program.py:156-240): returns (bgr uint8 [H,W,3], depth float32)."""
    R = pose_np.quat2mat(r) if r_type == "quat" else np.asarray(r)
    bgr, depth = _render(mesh, R, t, K, width, height, zNear, zFar, 2, light_position,
                         light_intensity, brightness_ratio)
    return bgr.astype(np.uint8), depth
