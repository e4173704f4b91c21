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
RC_ERR_INVALID_ARG, RC_ERR_CUDA, RC_ERR_UNALIGNED, RC_ERR_CAPACITY = 1, 2, 3, 4
# enums mirrored from include/rendercompare.h
RC_ZF_MASK, RC_ZF_IMAGE = 0, 1
RC_PRE_NONE, RC_PRE_THRESH02, RC_PRE_ADD_MEAN = 0, 1, 2
(RC_POST_NONE, RC_POST_ROUND, RC_POST_ROUND_M045, RC_POST_DIV_WX, RC_POST_MUL_WX,
 RC_POST_SUB_MEAN, RC_POST_SPOT_SUB_MEAN) = range(7)
RC_ROT_MODEL, RC_ROT_CAMERA, RC_ROT_CAMERA_NEW, RC_ROT_NAIVE = range(4)
RC_PM_L1, RC_PM_L2, RC_PM_SMOOTH_L1 = range(3)
RC_ZOOM_ACCUMULATE, RC_ZOOM_SRC_BROADCAST = 1, 2
RC_ZOOM_MAX_GROUPS = 12
RC_ZOOM_MAX_PLANES = 16
RC_ZF_WS_INTS = 10
RC_BOX_WS_INTS = 5

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
        ("flags", c_int32),
        ("mean", c_float * 4),
    ]


RC_STEP_MARKS = 6
RC_POSE_PART_POSE, RC_POSE_PART_LABELS = 1, 2
RC_STEP_PHASE_LOSS, RC_STEP_PHASE_FRAME, RC_STEP_PHASE_SAMPLE, RC_STEP_PHASE_ZT_BWD = 1, 2, 4, 8
_STEP_IN = ("image_observed", "image_rendered", "depth_gt_observed", "depth_rendered",
            "mask_observed", "mask_gt_observed", "src_pose", "KT", "Kinv", "K",
            "point_cloud_model", "point_cloud_weights", "point_cloud_observed", "rot_est",
            "trans_est", "trans_gt")
_STEP_OUT = ("flow", "valid", "zoom_factor", "bbox", "workspace", "zoom_mask_observed",
             "zoom_mask_gt_observed", "zoom_mask_rendered", "zoom_image_observed",
             "zoom_image_rendered", "zoom_flow", "zoom_flow_weights", "zoom_trans_gt", "loss_map",
             "loss_sum", "grad_rot", "grad_trans", "grad_zoom_trans")


class rc_compare_step_args(ctypes.Structure):
    """Mirror of `rc_compare_step_args` in include/rendercompare.h (field order matters)."""
    _fields_ = ([(n, c_void_p) for n in _STEP_IN] + [(n, c_void_p) for n in _STEP_OUT] + [
        ("pixel_means", c_float * 4),
        ("T_means", c_float * 3),
        ("T_stds", c_float * 3),
        ("norm_term", c_float),
        ("sl1_scalar", c_float),
        ("grad_scale", c_float),
        ("rot_coord", c_int32),
        ("loss_type", c_int32),
        ("zoom_grad", c_int32),
        ("batch_size", c_int32),
        ("height", c_int32),
        ("width", c_int32),
        ("num_points", c_int32),
        ("stream", c_void_p),
        ("marks", POINTER(c_void_p)),
        ("frame_renderer", c_void_p),
        ("observed_box", c_void_p),
        ("outputs_track_frames", c_int32),
        ("zoom_tile_state", c_void_p),
        ("loss_src_pose", c_void_p),
        ("depth_observed", c_void_p),
        ("zoom_depth_observed", c_void_p),
        ("zoom_depth_rendered", c_void_p),
        ("phases", c_int32),
    ])


RC_RENDER_TEXTURED, RC_RENDER_LAMBERT, RC_RENDER_LIT_LINEMOD = 0, 1, 2
RC_RENDER_MARKS = 5


class rc_mesh_desc(ctypes.Structure):
    _fields_ = [("vertices", c_void_p), ("uv", c_void_p), ("normals", c_void_p),
                ("triangles", c_void_p), ("texture", c_void_p), ("n_vertices", c_int32),
                ("n_triangles", c_int32), ("tex_h", c_int32), ("tex_w", c_int32)]


class rc_render_args(ctypes.Structure):
    _fields_ = [("cls_idx", c_void_p), ("poses", c_void_p), ("light_pos", c_void_p),
                ("light_int", c_void_p), ("depth", c_void_p), ("image_chw", c_void_p),
                ("bgr_hwc", c_void_p), ("mask", c_void_p), ("tri_id", c_void_p),
                ("status", c_void_p), ("K", c_float * 9), ("pixel_means", c_float * 3),
                ("znear", c_float), ("zfar", c_float), ("mask_thresh", c_float),
                ("brightness_ratio", c_float), ("mode", c_int32), ("batch_size", c_int32),
                ("incremental", c_int32), ("stream", c_void_p), ("marks", POINTER(c_void_p))]


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
    "rc_flow_host_release": (c_int, [c_int]),
    "rc_flow_dev": (c_int, [_P, _P, _P, _P, _P, _P, c_int, c_int, c_int, _P]),
    "rc_zoom_factor_dev": (c_int, [_P, _P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int,
                                   c_int, _P]),
    "rc_zoom_factor_noalloc_dev": (c_int, [_P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int,
                                           c_int, _P]),
    "rc_zoom_factor_noalloc_max_batch": (c_int, []),
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
    "rc_compare_step_dev": (c_int, [POINTER(rc_compare_step_args)]),
    "rc_event_create": (c_int, [POINTER(c_void_p)]),
    "rc_event_destroy": (c_int, [c_void_p]),
    "rc_event_record": (c_int, [c_void_p, c_void_p]),
    "rc_event_synchronize": (c_int, [c_void_p]),
    "rc_event_elapsed_ms": (c_int, [c_void_p, c_void_p, POINTER(c_float)]),
    "rc_renderer_create": (c_int, [POINTER(rc_mesh_desc), c_int, c_int, c_int, c_int, c_int,
                                   POINTER(c_void_p)]),
    "rc_renderer_destroy": (c_int, [c_void_p]),
    "rc_render_dev": (c_int, [c_void_p, POINTER(rc_render_args)]),
    "rc_renderer_note_replay": (c_int, [_P, c_int]),
    "rc_render_stats": (c_int, [c_void_p, c_int, c_void_p, POINTER(c_int64)]),
    "rc_pose_update_dev": (c_int, [_P, _P, _P, _P, _P, POINTER(c_float), POINTER(c_float), c_int,
                                   _P, _P, _P, _P, c_int, _P]),
    "rc_pose_delta_dev": (c_int, [_P, _P, POINTER(c_float), POINTER(c_float), c_int, _P, _P, c_int, _P]),
    "rc_standin_predict_dev": (c_int, [_P, _P, _P, _P, c_float, _P, _P, c_int, _P]),
    "rc_pose_update_standin_dev": (c_int, [_P, _P, _P, _P, _P, c_float, _P, _P, _P, _P,
                                           POINTER(c_float), POINTER(c_float), c_int, _P, _P, _P,
                                           _P, c_int, c_int, _P]),
    "rc_pm_loss_bwd_dev": (c_int, [_P, _P, _P, _P, _P, _P, POINTER(c_float), POINTER(c_float),
                                   c_int, c_int, c_float, c_float, c_float, _P, _P, c_int, c_int,
                                   _P]),
    "rc_mask_box_dev": (c_int, [_P, _P, _P, _P, c_int, c_int, c_int, _P]),
    "rc_pose_error_chunks": (c_int, [c_int]),
    "rc_pose_error_dev": (c_int, [_P, _P, _P, c_int, _P, _P, c_int, _P, _P, c_int, c_int, _P]),
    "rc_visib_mask_dev": (c_int, [_P, _P, _P, c_float, _P, c_int64, _P]),
    "rc_flow_epe_chunks": (c_int, []),
    "rc_flow_epe_dev": (c_int, [_P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, _P]),
    "rc_prep_image_dev": (c_int, [_P, _P, POINTER(ctypes.c_double), c_int, c_int, c_int, _P]),
    "rc_prep_depth_dev": (c_int, [_P, _P, _P, _P, c_float, c_int, c_int, c_int, _P]),
    "rc_prep_mask_dev": (c_int, [_P, _P, _P, c_int, c_int, c_int, _P]),
    "rc_mask_from_depth_dev": (c_int, [_P, _P, c_float, c_int, c_int64, _P]),
    "rc_mask_dilate_dev": (c_int, [_P, _P, _P, c_int, c_int, c_int, _P]),
    "rc_points_transform_dev": (c_int, [_P, _P, _P, c_int, c_int, _P]),
    "rc_calc_flow_blocks": (c_int, [c_int, c_int]),
    "rc_calc_flow_dev": (c_int, [_P, _P, _P, POINTER(ctypes.c_double), ctypes.c_double, c_int, _P,
                                 _P, _P, _P, _P, c_int, c_int, c_int, c_int, _P]),
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


class Event(object):
    """cudaEvent_t created through the C ABI (rc_event_*)."""

    def __init__(self):
        self.handle = c_void_p()
        check(load().rc_event_create(ctypes.byref(self.handle)), "rc_event_create")

    def record(self, stream=None):
        check(load().rc_event_record(self.handle, stream_ptr(stream)), "rc_event_record")

    def synchronize(self):
        check(load().rc_event_synchronize(self.handle), "rc_event_synchronize")

    def elapsed_ms(self, end) -> float:
        ms = c_float()
        check(load().rc_event_elapsed_ms(self.handle, end.handle, ctypes.byref(ms)),
              "rc_event_elapsed_ms")
        return float(ms.value)

    def __del__(self):
        try:
            if self.handle:
                load().rc_event_destroy(self.handle)
        except Exception:
            pass
