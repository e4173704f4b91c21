"""Fused on-device render-and-compare step.

One `CompareStep` owns every output buffer and workspace for a batch of B
pairs and issues, on the current CUDA stream and with no host synchronisation,
the launches that stand in for one pass of the reference's path between two
network evaluations:

    gpu_flow                     (lib/pair_matching/batch_updater_py_multi.py:331-352)
    ZoomMask                     (deepim/symbols/deepIM_flownet.py:595-606)
    ZoomImageWithFactor          (:614-623)
    ZoomFlow                     (:689-698)
    ZoomTrans                    (:659-665)
    Transform3D + PM loss        (:446-499)

ZoomMask's three masks, the six image channels, the two flow planes and the
two flow-weight planes share one sampling grid per pair, so they are ONE
launch of the resampler (13 planes); `np.tile(valid, [1,2,1,1])` is never
materialised (RC_ZOOM_SRC_BROADCAST).  ZoomTrans rides in the finalising block of the
bbox kernel.  Four launches per step in total, all issued by ONE C call
(rc_compare_step_dev).
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from . import _lib
from . import synth
from .operator_py import _common as C

# Algorithmic bytes per 640x480 pair: every input plane once, every
# output plane once.
BYTES_PER_PX_FLOW = 20
BYTES_PER_PX_ZOOM = 112
# the resampler launch alone: ZoomMask 12+12, ZoomImageWithFactor 24+24, ZoomFlow 16+16
# (the 8 B/px bbox pass over mask_gt_observed and mask_rendered is a separate kernel)
BYTES_PER_PX_ZOOM_SAMPLE = 104
BYTES_PER_PX_ZOOM_BBOX = 8


def algorithmic_bytes_per_pair(height, width, num_points, with_loss=True, with_flow=True,
                               with_zoom=True):
    px = height * width
    total = 0
    if with_flow:
        total += BYTES_PER_PX_FLOW * px
    if with_zoom:
        total += BYTES_PER_PX_ZOOM * px
    if with_loss:
        total += 4 * 3 * num_points * 4  # three [3,N] inputs + one [3,N] output
    return total


INPUT_KEYS = ("image_observed", "image_rendered", "depth_gt_observed", "depth_rendered",
              "mask_observed", "mask_gt_observed", "src_pose", "KT", "Kinv", "point_cloud_model",
              "point_cloud_weights", "point_cloud_observed", "rot_est", "trans_est", "trans_gt")


class CompareStep(object):
    """zoom-in + GT flow + point-matching loss for B pairs, entirely on one device."""

    def __init__(self, batch, device, height=synth.H, width=synth.W,
                 num_points=synth.NUM_3D_SAMPLE, K=synth.K_LINEMOD, pixel_means=synth.PIXEL_MEANS,
                 T_means=(0, 0, 0), T_stds=(1, 1, 1), rot_coord="CAMERA", loss_type="L1",
                 norm_term=synth.NORMALIZE_3D_POINT, want_loss_map=True):
        self.B, self.H, self.W, self.N = batch, height, width, num_points
        self.device = torch.device(device)
        self.K = np.asarray(K, np.float32)
        # the plugins receive PIXEL_MEANS and reverse it (zoom_image_with_factor.py:92-95)
        self.means = np.asarray(pixel_means, np.float32).reshape(-1)[::-1].copy()
        self.T_means, self.T_stds = _lib.f3(T_means), _lib.f3(T_stds)
        self.rot_coord = _lib.ROT_COORD[rot_coord.lower()]
        self.loss_type = _lib.PM_LOSS_TYPE[loss_type.lower()]
        self.norm_term = float(norm_term)
        f = dict(dtype=torch.float32, device=self.device)
        B, Hh, Ww, N = batch, height, width, num_points
        self.flow = torch.empty((B, 2, Hh, Ww), **f)
        self.valid = torch.empty((B, 1, Hh, Ww), **f)
        self.zoom_factor = torch.empty((B, 4), **f)
        self.bbox = torch.empty((B, 8), dtype=torch.int32, device=self.device)
        self.ws = torch.zeros((B, _lib.RC_ZF_WS_INTS), dtype=torch.int32, device=self.device)
        self.zoom_mask_observed = torch.empty((B, 1, Hh, Ww), **f)
        self.zoom_mask_gt_observed = torch.empty((B, 1, Hh, Ww), **f)
        self.zoom_mask_rendered = torch.empty((B, 1, Hh, Ww), **f)
        self.zoom_image_observed = torch.empty((B, 3, Hh, Ww), **f)
        self.zoom_image_rendered = torch.empty((B, 3, Hh, Ww), **f)
        self.zoom_flow = torch.empty((B, 2, Hh, Ww), **f)
        self.zoom_flow_weights = torch.empty((B, 2, Hh, Ww), **f)
        self.zoom_trans_gt = torch.empty((B, 3), **f)
        self.loss_map = torch.empty((B, 3, N), **f) if want_loss_map else None
        self.loss_sum = torch.empty((B,), **f)
        self.Kd = C.device_const(self.K, self.device)
        self.graph = None
        self.launches_per_step = 4  # flow, bbox+zoom factor(+ZoomTrans), resampler, pm loss
        self._lib = _lib.load()

    # ------------------------------------------------------------------ inputs
    def to_device(self, host, pinned=None, non_blocking=True):
        """Copy a host batch (numpy dict from synth.make_compare_batch, or pinned tensors)."""
        dev = {}
        for k in INPUT_KEYS:
            src = pinned[k] if pinned is not None else torch.from_numpy(np.ascontiguousarray(host[k]))
            dev[k] = src.to(self.device, non_blocking=non_blocking)
        return dev

    @staticmethod
    def input_bytes(host):
        return int(sum(np.asarray(host[k]).nbytes for k in INPUT_KEYS))

    # ------------------------------------------------------------------- step
    def bind(self, x):
        """Build the rc_compare_step_args for the input tensors in `x` (device, INPUT_KEYS).
        The struct holds raw pointers: keep `x` alive while it is in use."""
        a = _lib.rc_compare_step_args()
        for k in ("image_observed", "image_rendered", "depth_gt_observed", "depth_rendered",
                  "mask_observed", "mask_gt_observed", "src_pose", "KT", "Kinv",
                  "point_cloud_model", "point_cloud_weights", "point_cloud_observed", "rot_est",
                  "trans_est", "trans_gt"):
            t = x[k]
            C.check_tensor(k, t)
            if t.device != self.device:
                raise ValueError("%s is on %s, step is on %s" % (k, t.device, self.device))
            setattr(a, k, t.data_ptr())
        a.K = self.Kd.data_ptr()
        for name, t in (("flow", self.flow), ("valid", self.valid),
                        ("zoom_factor", self.zoom_factor), ("bbox", self.bbox),
                        ("workspace", self.ws), ("zoom_mask_observed", self.zoom_mask_observed),
                        ("zoom_mask_gt_observed", self.zoom_mask_gt_observed),
                        ("zoom_mask_rendered", self.zoom_mask_rendered),
                        ("zoom_image_observed", self.zoom_image_observed),
                        ("zoom_image_rendered", self.zoom_image_rendered),
                        ("zoom_flow", self.zoom_flow),
                        ("zoom_flow_weights", self.zoom_flow_weights),
                        ("zoom_trans_gt", self.zoom_trans_gt), ("loss_map", self.loss_map),
                        ("loss_sum", self.loss_sum)):
            setattr(a, name, t.data_ptr() if t is not None else None)
        for i in range(3):
            a.pixel_means[i] = float(self.means[i])
            a.T_means[i] = self.T_means[i]
            a.T_stds[i] = self.T_stds[i]
        a.norm_term = self.norm_term
        a.sl1_scalar = 1.0
        a.rot_coord = self.rot_coord
        a.loss_type = self.loss_type
        a.batch_size, a.height, a.width, a.num_points = self.B, self.H, self.W, self.N
        a.marks = None
        return a

    def run(self, x, stream=None, marks=None, args=None):
        """One step: a single C call that enqueues the four kernels on the current stream.
        `marks`: optional list of RC_STEP_MARKS _lib.Event (or None entries) recorded around
        the kernels.  `args`: a struct from bind(x) to skip re-binding."""
        a = self.bind(x) if args is None else args
        a.stream = _lib.stream_ptr(stream)
        if marks is not None:
            arr = (ctypes.c_void_p * _lib.RC_STEP_MARKS)()
            for i, m in enumerate(marks):
                arr[i] = m.handle if m is not None else None
            a.marks = ctypes.cast(arr, ctypes.POINTER(ctypes.c_void_p))
        else:
            a.marks = None
        with torch.cuda.device(self.device):
            st = self._lib.rc_compare_step_dev(ctypes.byref(a))
        _lib.check(st, "rc_compare_step_dev")

    # ------------------------------------------------------------- CUDA graph
    def capture(self, x):
        """Capture run(x) into a CUDA graph (inputs are read from the tensors in `x` in place)."""
        self.run(x)  # warm-up outside capture
        torch.cuda.synchronize(self.device)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self.run(x)
        self.graph = g
        return g

    def outputs(self):
        return {
            "flow": self.flow, "flow_weights_valid": self.valid, "zoom_factor": self.zoom_factor,
            "bbox": self.bbox, "zoom_mask_observed": self.zoom_mask_observed,
            "zoom_mask_gt_observed": self.zoom_mask_gt_observed,
            "zoom_mask_rendered": self.zoom_mask_rendered,
            "zoom_image_observed": self.zoom_image_observed,
            "zoom_image_rendered": self.zoom_image_rendered, "zoom_flow": self.zoom_flow,
            "zoom_flow_weights": self.zoom_flow_weights, "zoom_trans_gt": self.zoom_trans_gt,
            "point_matching_loss": self.loss_map, "point_matching_loss_sum": self.loss_sum,
        }
