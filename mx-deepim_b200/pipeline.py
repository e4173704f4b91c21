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
                 norm_term=synth.NORMALIZE_3D_POINT, want_loss_map=True, backward=False,
                 lw_pm=1.0, zoom_grad=False, input_depth=False):
        """input_depth=True (cfg.network.INPUT_DEPTH, deepIM_flownet.py:669-681 / :827-835): the
        step also zooms `depth_observed` and `depth_rendered` (ZoomDepth) as two more planes of
        its one resampler launch; bind() then needs x["depth_observed"].

        backward=True adds the training backward leg to every step: d loss / d rot_est,
        d loss / d trans_est (Transform3D.backward under MakeLoss(grad_scale=LW_PM/N),
        transform3d.py:120-327, deepIM_flownet.py:493-498) and the gradient the network's
        zoomed translation head receives (invZoomTrans.backward, zoom_trans.py:50-76,
        b_zoom_grad=`zoom_grad`).  The zoom ops' own backwards are constant zeros
        (zoom_mask.py:136-140 ...) and are not materialised."""
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
        self.input_depth = bool(input_depth)
        self.zoom_depth_observed = torch.empty((B, 1, Hh, Ww), **f) if input_depth else None
        self.zoom_depth_rendered = torch.empty((B, 1, Hh, Ww), **f) if input_depth else None
        self.loss_map = torch.empty((B, 3, N), **f) if want_loss_map else None
        self.loss_sum = torch.empty((B,), **f)
        self.backward = bool(backward)
        self.grad_scale = float(lw_pm) / float(max(N, 1))
        self.zoom_grad = bool(zoom_grad)
        self.grad_rot = torch.empty((B, 4), **f) if backward else None
        self.grad_trans = torch.empty((B, 3), **f) if backward else None
        self.grad_zoom_trans = torch.empty((B, 3), **f) if backward else None
        # one flag per 32x8 block of resampler outputs (sparse frame step, see RefineLoop)
        self.zoom_tile_state = torch.zeros((B, ((Hh + 7) // 8) * ((Ww + 127) // 128) * 4),
                                           dtype=torch.uint8, device=self.device)
        self.Kd = C.device_const(self.K, self.device)
        self.graph = None
        # flow, bbox+zoom factor(+ZoomTrans), resampler, pm loss (+ loss bwd, ZoomTrans bwd)
        self.launches_per_step = 6 if backward else 4
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
    def bind(self, x, test_phase=False):
        """Build the rc_compare_step_args for the input tensors in `x` (device, INPUT_KEYS).
        The struct holds raw pointers: keep `x` alive while it is in use.

        test_phase=True mirrors the test graph (deepim/symbols/deepIM_flownet.py:778-838):
        there is no ground truth, so `mask_gt_observed` IS `mask_observed` (:781), no GT flow
        and no point-matching loss are computed; only image_observed, image_rendered,
        depth_rendered, mask_observed and src_pose are read."""
        a = _lib.rc_compare_step_args()
        keys = ("image_observed", "image_rendered", "depth_gt_observed", "depth_rendered",
                "mask_observed", "mask_gt_observed", "src_pose", "KT", "Kinv",
                "point_cloud_model", "point_cloud_weights", "point_cloud_observed", "rot_est",
                "trans_est", "trans_gt")
        if test_phase:
            keys = ("image_observed", "image_rendered", "depth_rendered", "mask_observed",
                    "src_pose")
        for k in keys:
            t = x[k]
            C.check_tensor(k, t)
            if t.device != self.device:
                raise ValueError("%s is on %s, step is on %s" % (k, t.device, self.device))
            setattr(a, k, t.data_ptr())
        a.K = self.Kd.data_ptr()
        lsp = x.get("loss_src_pose")
        if lsp is not None and not test_phase:
            C.check_tensor("loss_src_pose", lsp)
            a.loss_src_pose = lsp.data_ptr()
        if test_phase:
            a.mask_gt_observed = x["mask_observed"].data_ptr()
        if self.input_depth:
            t = x["depth_observed"]
            C.check_tensor("depth_observed", t)
            if t.device != self.device:
                raise ValueError("depth_observed is on %s, step is on %s" % (t.device, self.device))
            a.depth_observed = t.data_ptr()
            a.zoom_depth_observed = self.zoom_depth_observed.data_ptr()
            a.zoom_depth_rendered = self.zoom_depth_rendered.data_ptr()
        skip = ("flow", "valid", "zoom_flow", "zoom_flow_weights", "zoom_trans_gt", "loss_map",
                "loss_sum", "zoom_mask_gt_observed", "grad_rot", "grad_trans",
                "grad_zoom_trans") if test_phase else ()
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
                        ("loss_sum", self.loss_sum), ("grad_rot", self.grad_rot),
                        ("grad_trans", self.grad_trans),
                        ("grad_zoom_trans", self.grad_zoom_trans)):
            setattr(a, name, t.data_ptr() if (t is not None and name not in skip) else None)
        for i in range(3):
            a.pixel_means[i] = float(self.means[i])
            a.T_means[i] = self.T_means[i]
            a.T_stds[i] = self.T_stds[i]
        a.norm_term = self.norm_term
        a.sl1_scalar = 1.0
        a.grad_scale = self.grad_scale
        a.zoom_grad = 1 if self.zoom_grad else 0
        a.rot_coord = self.rot_coord
        a.loss_type = self.loss_type
        a.batch_size, a.height, a.width, a.num_points = self.B, self.H, self.W, self.N
        if test_phase:
            a.num_points = 0
        a.marks = None
        return a

    def run(self, x, stream=None, marks=None, args=None):
        """One step: a single C call that enqueues the four kernels on the current stream.
        `marks`: optional list of RC_STEP_MARKS _lib.Event (or None entries) recorded around
        the kernels.  `args`: a struct from bind(x) to skip re-binding."""
        a = self.bind(x) if args is None else args
        with torch.cuda.device(self.device):  # the CURRENT stream of the step's device
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
            "grad_rot": self.grad_rot, "grad_trans": self.grad_trans,
            "grad_zoom_trans": self.grad_zoom_trans,
            "zoom_depth_observed": self.zoom_depth_observed,
            "zoom_depth_rendered": self.zoom_depth_rendered,
        }


PACKED_KEYS = ("image_observed", "image_rendered", "depth_gt_observed", "depth_rendered",
               "mask_observed", "mask_gt_observed")


class HostFeed(object):
    """Feeds a CompareStep from pinned host batches with the upload of batch i+1 overlapped
    with the step of batch i: `depth` device staging sets, uploads on a private copy stream,
    the step (and the storage-format decode, if `packed`) on the caller's current stream.

    packed=False: every INPUT_KEYS entry is float32 as the reference's data loader hands it to
    the executor.  packed=True: the six image-sized inputs arrive in the dataset's storage
    formats (uint8 BGR HWC images, 16-bit depth in 1/depth_factor m, uint8 instance masks) and
    are decoded on the device (prep.py: image.transform, DEPTH_FACTOR, mask == idx) into the
    float32 staging set; the small inputs stay float32.

    submit(host) returns the staging-set index i; after a synchronize (or `done[i]`) the
    step's loss sums / zoom factors for that batch are in `loss[i]` / `zoom_factor[i]` (pinned).
    """

    def __init__(self, step, example, packed=False, depth_factor=1000.0, pixel_means_bgr=None,
                 depth=2):
        from . import prep

        self.prep = prep
        self.step, self.packed, self.depth = step, bool(packed), int(depth)
        self.depth_factor = float(depth_factor)
        self.means_bgr = np.asarray(
            synth.PIXEL_MEANS if pixel_means_bgr is None else pixel_means_bgr, np.float64).reshape(-1)
        dev, B = step.device, step.B
        f = dict(dtype=torch.float32, device=dev)
        self.sets, self.raw, self.args = [], [], []
        for _ in range(self.depth):
            st = {}
            for k in INPUT_KEYS:
                e = example[k]
                if self.packed and k in PACKED_KEYS:
                    ch = 3 if k.startswith("image") else 1
                    st[k] = torch.empty((B, ch, step.H, step.W), **f)
                else:
                    st[k] = torch.empty(tuple(e.shape), **f)
            self.sets.append(st)
            self.raw.append({k: torch.empty_like(example[k], device=dev) for k in PACKED_KEYS}
                            if self.packed else None)
            self.args.append(step.bind(st))
        self.one = torch.ones((B,), dtype=torch.int32, device=dev)
        self.copy_stream = torch.cuda.Stream(device=dev)
        self.uploaded = [torch.cuda.Event() for _ in range(self.depth)]
        self.done = [torch.cuda.Event() for _ in range(self.depth)]
        self.loss = torch.empty((self.depth, B), dtype=torch.float32).pin_memory()
        self.zoom_factor = torch.empty((self.depth, B, 4), dtype=torch.float32).pin_memory()
        self.n = 0
        self.h2d_bytes = int(sum(example[k].numel() * example[k].element_size() for k in INPUT_KEYS))
        self.d2h_bytes = int(B * 4 + B * 16)
        self.launches_per_submit = step.launches_per_step + (6 if self.packed else 0)

    def submit(self, host):
        i = self.n % self.depth
        compute = torch.cuda.current_stream(self.step.device)
        st, raw = self.sets[i], self.raw[i]
        with torch.cuda.stream(self.copy_stream):
            if self.n >= self.depth:  # the set's previous batch has been consumed
                self.copy_stream.wait_event(self.done[i])
            for k in INPUT_KEYS:
                dst = raw[k] if (self.packed and k in PACKED_KEYS) else st[k]
                dst.copy_(host[k], non_blocking=True)
            self.uploaded[i].record(self.copy_stream)
        compute.wait_event(self.uploaded[i])
        if self.packed:
            p = self.prep
            p.transform_batch(raw["image_observed"], self.means_bgr, out=st["image_observed"])
            p.transform_batch(raw["image_rendered"], self.means_bgr, out=st["image_rendered"])
            p.decode_depth(raw["depth_gt_observed"], self.depth_factor, out=st["depth_gt_observed"])
            p.decode_depth(raw["depth_rendered"], self.depth_factor, out=st["depth_rendered"])
            p.decode_mask(raw["mask_observed"], self.one, out=st["mask_observed"])
            p.decode_mask(raw["mask_gt_observed"], self.one, out=st["mask_gt_observed"])
        self.step.run(None, args=self.args[i])
        self.loss[i].copy_(self.step.loss_sum, non_blocking=True)
        self.zoom_factor[i].copy_(self.step.zoom_factor, non_blocking=True)
        self.done[i].record(compute)
        self.n += 1
        return i


def mask_box(mask, out=None, bbox=None, workspace=None, stream=None):
    """Test-time mask policy "box_rendered" / "box_observed" of update_data_batch
    (/root/reference/lib/pair_matching/data_pair.py:101-131): the bounding rectangle of the
    non-zero pixels of `mask` [B,1,H,W] (exclusive slice end, as the reference writes it).
    Returns (out [B,1,H,W], bbox [B,4] int32 = x_start, x_end, y_start, y_end); an empty mask
    gives x_start > x_end and an all-zero `out` (the reference raises there -- check `bbox`
    on the host if that matters).  No synchronisation."""
    C.check_tensor("mask", mask)
    B, _, H, W = mask.shape
    dev = mask.device
    if out is None:
        out = torch.empty_like(mask)
    if bbox is None:
        bbox = torch.empty((B, 4), dtype=torch.int32, device=dev)
    if workspace is None:
        workspace = torch.zeros((B, _lib.RC_BOX_WS_INTS), dtype=torch.int32, device=dev)
    if out.data_ptr() == mask.data_ptr():
        raise ValueError("out must not alias mask")
    with torch.cuda.device(dev):
        st = _lib.load().rc_mask_box_dev(mask.data_ptr(), out.data_ptr(), bbox.data_ptr(),
                                         workspace.data_ptr(), B, H, W, _lib.stream_ptr(stream))
    _lib.check(st, "rc_mask_box_dev")
    return out, bbox


def pose_update(src_pose, rot_est, trans_est, tgt_pose, K, T_means=(0, 0, 0), T_stds=(1, 1, 1),
                rot_coord="CAMERA", out=None, stream=None):
    """Batched pose block of the batch updater (batch_updater_py_multi.py:211-312) on device.
    Returns (refined_pose [B,3,4], rot_res [B,4], trans_res [B,3], KT [B,3,4])."""
    for n, t in (("src_pose", src_pose), ("rot_est", rot_est), ("trans_est", trans_est),
                 ("tgt_pose", tgt_pose)):
        C.check_tensor(n, t)
    B = src_pose.shape[0]
    dev = src_pose.device
    if out is None:
        f = dict(dtype=torch.float32, device=dev)
        out = (torch.empty((B, 3, 4), **f), torch.empty((B, 4), **f), torch.empty((B, 3), **f),
               torch.empty((B, 3, 4), **f))
    if out[0].data_ptr() == src_pose.data_ptr():
        raise ValueError("refined_pose must not alias src_pose")
    Kd = C.device_const(K, dev)
    with torch.cuda.device(dev):
        st = _lib.load().rc_pose_update_dev(
            src_pose.data_ptr(), rot_est.data_ptr(), trans_est.data_ptr(), tgt_pose.data_ptr(),
            Kd.data_ptr(), _lib.f3(T_means), _lib.f3(T_stds), _lib.ROT_COORD[rot_coord.lower()],
            out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(), out[3].data_ptr(), B,
            _lib.stream_ptr(stream))
    _lib.check(st, "rc_pose_update_dev")
    return out


def pose_delta(src_pose, tgt_pose, T_means=(0, 0, 0), T_stds=(1, 1, 1), rot_coord="CAMERA",
               out=None, stream=None):
    """Batched calc_RT_delta(src, tgt, rot_type="QUAT") (RT_transform.py:16-48) on device: the
    iteration-0 regression labels.  Returns (rot [B,4], trans [B,3])."""
    C.check_tensor("src_pose", src_pose)
    C.check_tensor("tgt_pose", tgt_pose)
    B = src_pose.shape[0]
    dev = src_pose.device
    if out is None:
        f = dict(dtype=torch.float32, device=dev)
        out = (torch.empty((B, 4), **f), torch.empty((B, 3), **f))
    with torch.cuda.device(dev):
        st = _lib.load().rc_pose_delta_dev(
            src_pose.data_ptr(), tgt_pose.data_ptr(), _lib.f3(T_means), _lib.f3(T_stds),
            _lib.ROT_COORD[rot_coord.lower()], out[0].data_ptr(), out[1].data_ptr(), B,
            _lib.stream_ptr(stream))
    _lib.check(st, "rc_pose_delta_dev")
    return out


class RefineLoop(object):
    """On-device replacement of the refine loop's data path
    (/root/reference/deepim/core/module.py:1205-1213 calling
    /root/reference/lib/pair_matching/batch_updater_py_multi.py:100-387 between network passes):

        per iteration:  network stand-in -> pose update -> rasterise at the refined pose
                        (+ fused image post / mask) -> GT flow -> zoom (13 planes) -> PM loss

    Everything stays on the device; one iteration is 1 (pose) + 1 memset + 7 (render) +
    4 (compare step) launches plus the stand-in's few elementwise torch kernels, and no host
    synchronisation.  The CNN is out of scope (north_star): `predict` is a pluggable callback
    (labels -> predictions); the default is the damped ground-truth update of SURVEY.md 8(d).
    """

    def __init__(self, meshes, batch, device, K=synth.K_LINEMOD, pixel_means=synth.PIXEL_MEANS,
                 num_points=synth.NUM_3D_SAMPLE, damping=0.5, predict=None, seed=0,
                 backward=False, lw_pm=1.0, sparse=True, side_stream=True):
        """side_stream=True: the point-matching loss (+ backward) of an iteration is issued on a
        second stream right after the pose update -- it reads poses and points only, so it runs
        beside the rasteriser's latency-bound passes instead of in front of the compare step --
        and the ZoomTrans backward beside the resampler (rc_compare_step_args.phases).  The
        iteration joins the side stream before it ends; inside a captured loop the fork / join
        become edges of the graph.  Results are identical (same kernels, same inputs).

        sparse=True: the compare step takes the rendered crop box and the GT flow from the
        16x16 tiles the renderer reports as covered (rc_compare_step_args.frame_renderer,
        csrc/frame.cu) instead of streaming the ~90 % background of both planes; results are
        identical (tests/test_refine_gpu.py)."""
        from .render import DeviceRenderer

        self.sparse = bool(sparse)
        self.B = batch
        self.device = torch.device(device)
        self.K = np.asarray(K, np.float32)
        self.renderer = DeviceRenderer(meshes, self.K, synth.W, synth.H, synth.ZNEAR, synth.ZFAR,
                                       max_batch=batch, device=self.device)
        self.step = CompareStep(batch, self.device, K=self.K, pixel_means=pixel_means,
                                num_points=num_points, backward=backward, lw_pm=lw_pm)
        # the updater subtracts PIXEL_MEANS[[2,1,0]] from the RGB image (:24-25, :279-284)
        self.render_means = np.asarray(pixel_means, np.float32)[[2, 1, 0]].copy()
        f = dict(dtype=torch.float32, device=self.device)
        B = batch
        self.pose = [torch.empty((B, 3, 4), **f), torch.empty((B, 3, 4), **f)]
        self.rot_gt = torch.empty((B, 4), **f)
        self.trans_gt = torch.empty((B, 3), **f)
        self.KT = torch.empty((B, 3, 4), **f)
        self.depth_rendered = torch.empty((B, 1, synth.H, synth.W), **f)
        self.image_rendered = torch.empty((B, 3, synth.H, synth.W), **f)
        self.mask_rendered = torch.empty((B, 1, synth.H, synth.W), **f)
        self.rot_est = torch.empty((B, 4), **f)
        self.trans_est = torch.empty((B, 3), **f)
        # the batch's observed side lives in PERSISTENT tensors (set_batch copies into them), so
        # bound argument structs and captured graphs stay valid from batch to batch
        Hh, Ww, N = synth.H, synth.W, num_points
        self.obs = {"image_observed": torch.zeros((B, 3, Hh, Ww), **f),
                    "depth_gt_observed": torch.zeros((B, 1, Hh, Ww), **f),
                    "mask_observed": torch.zeros((B, 1, Hh, Ww), **f),
                    "mask_gt_observed": torch.zeros((B, 1, Hh, Ww), **f),
                    "tgt_pose": torch.zeros((B, 3, 4), **f),
                    "point_cloud_model": torch.zeros((B, 3, N), **f),
                    "point_cloud_weights": torch.zeros((B, 3, N), **f),
                    "point_cloud_observed": torch.zeros((B, 3, N), **f),
                    "Kinv": torch.zeros((3, 3), **f)}
        self.cls = torch.zeros((B,), dtype=torch.int32, device=self.device)
        self.pose0 = torch.zeros((B, 3, 4), **f)   # the batch's initial source poses
        self.rot0 = torch.zeros((B, 4), **f)       # iteration-0 labels (calc_RT_delta(src, tgt))
        self.trans0 = torch.zeros((B, 3), **f)
        self.observed_box = torch.zeros((B, 4), dtype=torch.int32, device=self.device)
        self._zf_scratch = torch.empty((B, 4), **f)
        self._box8 = torch.empty((B, 8), dtype=torch.int32, device=self.device)
        self._have_batch = False
        self.damping = float(damping)
        self.predict = predict  # None: the damped-GT stand-in, fused into the pose update's launch
        self.side = torch.cuda.Stream(device=self.device) if side_stream else None
        self._ev_fork, self._ev_zf, self._ev_join, self._ev_labels = (torch.cuda.Event() for _ in range(4))
        g = torch.Generator(device="cpu").manual_seed(int(seed))
        self.noise_q = (0.01 * torch.randn((8, B, 4), generator=g)).to(self.device)
        self.noise_t = (0.005 * torch.randn((8, B, 3), generator=g)).to(self.device)
        self.identity_q = torch.tensor([1.0, 0.0, 0.0, 0.0], **f)
        self.cur = 0
        self.graph = None
        self._flow_tracks = False
        self.bound = [None, None]
        for par in (0, 1):
            x = dict(self.obs)
            # flow / zoom read the refined pose pose[par]; the loss (and its backward) belongs to
            # the pass that produced rot_est / trans_est, i.e. to the pose BEFORE the update
            # (deepIM_flownet.py:446-499 vs batch_updater_py_multi.py:211-312)
            x.update(image_rendered=self.image_rendered, depth_rendered=self.depth_rendered,
                     src_pose=self.pose[par], loss_src_pose=self.pose[1 - par], KT=self.KT,
                     rot_est=self.rot_est, trans_est=self.trans_est, trans_gt=self.trans_gt)
            self.bound[par] = self.step.bind(x)
            self._bind_frame(self.bound[par])
        # pose update (+ fused stand-in) + 7 render kernels (clear, vertex, setup, scan, fill,
        # cover, resolve) + step
        # (with the side stream the pose update is two launches: pose | labels + KT)
        # (sparse, small batches: the frame pass's last block computes the zoom factors -- one
        # launch less than CompareStep counts)
        self.launches_per_iteration = (2 if (side_stream or predict is not None) else 1) + 7 + \
            self.step.launches_per_step - (1 if (sparse and batch <= 64) else 0)

    def set_batch(self, host):
        """Observed side of the batch + initial poses.  `host`: numpy dict with image_observed,
        depth_gt_observed, mask_observed, mask_gt_observed, src_pose, tgt_pose, class_index,
        point_cloud_model / weights / observed, Kinv (float32 blobs as the reference's loader
        hands them over).  Copies into the loop's persistent tensors; captured graphs stay valid."""
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a))
        for k in self.obs:
            self.obs[k].copy_(t(host[k]))
        self.cls.copy_(t(np.asarray(host["class_index"], np.int32)))
        self.pose0.copy_(t(host["src_pose"]))
        self.finish_batch()

    def finish_batch(self, stream=None):
        """After the observed side, `cls` and `pose0` have been written (set_batch, or a host
        feed's on-device decode): iteration-0 labels, the observed crop box, loop restart.
        No synchronisation."""
        # iteration-0 regression labels: calc_RT_delta(src, tgt) (the loader's job in the reference)
        pose_delta(self.pose0, self.obs["tgt_pose"], out=(self.rot0, self.trans0), stream=stream)
        if self.sparse:
            # the observed box is a property of the batch: one dense pass now instead of one per
            # iteration (obs == rend == mask_gt_observed; entries 0..3 are the "> 0.3" box)
            C.zoom_factor(self.obs["mask_gt_observed"], self.obs["mask_gt_observed"], self.pose0,
                          self.K, None, self._zf_scratch, _lib.RC_ZF_MASK, bbox=self._box8,
                          stream=stream)
            self.observed_box.copy_(self._box8[:, :4])
        self._have_batch = True
        self.set_batch_poses()

    def _bind_frame(self, a):
        if self.sparse:
            a.frame_renderer = self.renderer.handle
            a.observed_box = self.observed_box.data_ptr()
            a.zoom_tile_state = self.step.zoom_tile_state.data_ptr()
            a.outputs_track_frames = 0

    def set_batch_poses(self, host=None):
        """Restart the loop from the batch's initial source poses (same observed side).
        `host` is accepted for compatibility and ignored: the poses are the ones set_batch got."""
        if not self._have_batch:
            raise RuntimeError("RefineLoop: set_batch() first")
        self.pose[0].copy_(self.pose0)
        self.rot_gt.copy_(self.rot0)
        self.trans_gt.copy_(self.trans0)
        self.cur = 0

    def _pose_update(self, it, nxt, parts=3, stream=None):
        """pose[cur] (+ prediction) -> pose[nxt] (parts & 1), new labels + KT (parts & 2).  With
        the default stand-in the prediction is made at the head of the pose part's launch; the
        labels part reads the prediction back from rot_est / trans_est."""
        out = (self.pose[nxt], self.rot_gt, self.trans_gt, self.KT)
        if self.predict is not None:
            self.predict(it, self.rot_gt, self.trans_gt)
            pose_update(self.pose[self.cur], self.rot_est, self.trans_est, self.obs["tgt_pose"],
                        self.K, out=out)
            return
        standin = bool(parts & _lib.RC_POSE_PART_POSE)
        with torch.cuda.device(self.device):
            st = _lib.load().rc_pose_update_standin_dev(
                self.pose[self.cur].data_ptr(),
                self.rot_gt.data_ptr() if standin else None,
                self.trans_gt.data_ptr() if standin else None,
                self.noise_q[it % 8].data_ptr() if standin else None,
                self.noise_t[it % 8].data_ptr() if standin else None, self.damping,
                self.rot_est.data_ptr(), self.trans_est.data_ptr(), self.obs["tgt_pose"].data_ptr(),
                C.device_const(self.K, self.device).data_ptr(), _lib.f3((0, 0, 0)),
                _lib.f3((1, 1, 1)), _lib.ROT_COORD["camera"], out[0].data_ptr(),
                out[1].data_ptr(), out[2].data_ptr(), out[3].data_ptr(), int(parts), self.B,
                _lib.stream_ptr(stream))
        _lib.check(st, "rc_pose_update_standin_dev")

    def iterate(self, it, marks=None, render_marks=None):
        """One refinement iteration on the current stream (no host synchronisation)."""
        nxt = 1 - self.cur
        if self.renderer.serial != getattr(self, "_my_serial", -1):
            # somebody else rendered through this renderer since our last frame: its item lists no
            # longer describe what the flow planes / zoomed outputs hold -> one dense step
            self._flow_tracks = False
        a = self.bound[nxt]
        # the flow / bbox / loss marks serialise the step (clean per-kernel times): one call, one
        # stream; the resampler's marks (2, 3) do not and go with the SAMPLE phase
        serial = marks is not None and any(marks[i] is not None for i in (0, 1, 4, 5))
        split = self.side is not None and not serial and bool(a.loss_sum or a.grad_rot)
        # the rasteriser waits for the refined pose only: the new labels and KT (mat2quat's
        # float64 eigen-solve, most of the pose kernel's latency) go to the side stream as well
        split_pose = split and self.predict is None
        self._pose_update(it, nxt, parts=_lib.RC_POSE_PART_POSE if split_pose else 3)
        if split:
            main = torch.cuda.current_stream(self.device)
            self._ev_fork.record(main)
            self.side.wait_event(self._ev_fork)
            if split_pose:
                self._pose_update(it, nxt, parts=_lib.RC_POSE_PART_LABELS, stream=self.side)
                self._ev_labels.record(self.side)
            a.phases = _lib.RC_STEP_PHASE_LOSS
            self.step.run(None, args=a, stream=self.side)
        self.renderer.render_batch(self.cls, self.pose[nxt], pixel_means=self.render_means,
                                   want=("depth", "image", "mask"),
                                   out={"depth": self.depth_rendered, "image": self.image_rendered,
                                        "mask": self.mask_rendered}, marks=render_marks,
                                   incremental=True)  # the loop owns these three buffers
        self._my_serial = self.renderer.serial
        self.cur = nxt
        if self.sparse:
            a.outputs_track_frames = 1 if self._flow_tracks else 0
        if split:
            if split_pose:
                main.wait_event(self._ev_labels)  # KT (GT flow) and trans_gt (ZoomTrans)
            a.phases = _lib.RC_STEP_PHASE_FRAME
            self.step.run(None, args=a)
            self._ev_zf.record(main)
            a.phases = _lib.RC_STEP_PHASE_SAMPLE
            self.step.run(None, args=a, marks=marks)
            self.side.wait_event(self._ev_zf)
            a.phases = _lib.RC_STEP_PHASE_ZT_BWD
            self.step.run(None, args=a, stream=self.side)
            self._ev_join.record(self.side)
            main.wait_event(self._ev_join)
            a.phases = 0
        else:
            a.phases = 0
            self.step.run(None, args=a, marks=marks)
        self._flow_tracks = True  # one step per render: the flow planes follow the frames

    def run(self, iterations=4, marks=None):
        """`marks`: optional per-iteration lists of RC_STEP_MARKS events (see CompareStep.run)."""
        for it in range(iterations):
            self.iterate(it, marks=marks[it] if marks is not None else None)

    def check_status(self):
        """Synchronise and raise if any render of this loop overflowed the rasteriser's tile
        lists (DeviceRenderer.check_status).  The loop itself never synchronises; call this at
        the end of a loop / epoch, before trusting its outputs."""
        self.renderer.check_status()

    def capture(self, host=None, iterations=4, marks=None):
        """Capture one whole loop -- restart from the batch's initial poses, then `iterations`
        refinement iterations (stand-in, pose update, render, compare step: ~14 launches and 2
        memsets each) -- into a CUDA graph; `self.replay()` then costs one launch call from the
        host, which is what matters once a GPU's share of the batch is small (8 GPUs x 32
        pairs).  The programmatic dependent launches inside the render and the step are kept
        as programmatic edges of the graph.

        The renderer's incremental background ping-pongs between two item lists per render, and
        a replay has to find the previous loop's last frame in the half it was captured to read.
        An even `iterations` returns to the same half, so one graph serves every replay; for an
        odd count TWO graphs are captured (one per starting half) and replay() alternates
        between them.  replay() also notices when anybody rendered through this loop's renderer
        since the last replay (DeviceRenderer.serial) -- the captured pointers would then read
        another frame's item list -- and re-captures instead of replaying stale state."""
        if iterations < 1:
            raise ValueError("capture() needs at least one iteration")
        self.set_batch_poses(host)
        self.run(iterations)  # warm-up; also establishes the incremental-background state
        torch.cuda.synchronize(self.device)
        self.graphs = []
        for _ in range(1 if iterations % 2 == 0 else 2):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self.set_batch_poses(host)
                # marks: event-record NODES inside the graph (rc::record_mark), re-recorded by
                # every replay -- kernel timing inside a replayed loop
                self.run(iterations, marks=marks)
            self.graphs.append(g)
            # capturing ran the HOST side of every render (the handle's ping-pong parity moved)
            # but not the kernels: execute the loop once so that device and host agree again
            g.replay()
        self._graph_turn = 0
        self.graph = self.graphs[0]
        self.graph_iterations = iterations
        self._graph_marks = marks
        self._graph_serial = self.renderer.serial
        return self.graph

    def replay(self):
        """Replay the captured loop (capture() first).  Safe against an eager render through the
        same renderer in between (re-captures) and for odd iteration counts (alternates)."""
        if getattr(self, "graph", None) is None:
            raise RuntimeError("RefineLoop.replay(): no captured loop (capture() first, and again "
                               "after set_batch)")
        if self.renderer.serial != self._graph_serial:
            self.capture(None, self.graph_iterations, marks=self._graph_marks)
        g = self.graphs[self._graph_turn % len(self.graphs)]
        self._graph_turn += 1
        g.replay()
        if self.graph_iterations % 2:
            self.renderer.note_replay(self.graph_iterations)

    # ------------------------------------------------------------------ test phase
    def iterate_test(self, rot_est, trans_est, update_mask="box_rendered"):
        """One TEST-time refinement iteration (deepim/core/tester.py:520-594 +
        lib/pair_matching/data_pair.py:75-138) on the current stream, no synchronisation:

            pose_rendered_update = RT_transform(pose_rendered, r, t)     (tester.py:520-530)
            render at the new pose; mask_rendered = depth > 0.2          (:557-577)
            mask_observed <- policy(update_mask)                         (:578-589, data_pair)
            zoom-in of masks + images through the test graph             (deepIM_flownet :778-838)

        update_mask: "init" keeps mask_observed; "box_rendered" replaces it with the bounding
        rectangle of the PREVIOUS iteration's rendered mask (tester.py:583-585 reads the batch's
        current mask_rendered blob before it is overwritten)."""
        if update_mask not in ("init", "box_rendered"):
            raise Exception("Unknown UPDATE_MASK type: {}".format(update_mask))  # tester.py:588
        nxt = 1 - self.cur
        if update_mask == "box_rendered":
            if not hasattr(self, "_box_ws"):
                self._box_ws = torch.zeros((self.B, _lib.RC_BOX_WS_INTS), dtype=torch.int32,
                                           device=self.device)
                self._box_bbox = torch.empty((self.B, 4), dtype=torch.int32, device=self.device)
                self.mask_observed_test = torch.empty_like(self.mask_rendered)
            mask_box(self.mask_rendered, out=self.mask_observed_test, bbox=self._box_bbox,
                     workspace=self._box_ws)
        pose_update(self.pose[self.cur], rot_est, trans_est, self.pose[self.cur], self.K,
                    out=(self.pose[nxt], self._scratch_q(), self._scratch_t(), self.KT))
        self.renderer.render_batch(self.cls, self.pose[nxt], pixel_means=self.render_means,
                                   want=("depth", "image", "mask"),
                                   out={"depth": self.depth_rendered, "image": self.image_rendered,
                                        "mask": self.mask_rendered}, incremental=True)
        self._my_serial = self.renderer.serial
        self.cur = nxt
        mo = self.mask_observed_test if update_mask == "box_rendered" else self.obs["mask_observed"]
        x = {"image_observed": self.obs["image_observed"], "image_rendered": self.image_rendered,
             "depth_rendered": self.depth_rendered, "mask_observed": mo,
             "src_pose": self.pose[nxt]}
        # dense step: in the test graph the "observed" box is that of mask_observed, which the
        # mask policy changes every iteration, so there is no per-batch box to reuse
        self.step.run(None, args=self.step.bind(x, test_phase=True))
        self._flow_tracks = False  # this frame was rendered without a flow pass

    def _scratch_q(self):
        if not hasattr(self, "_sq"):
            self._sq = torch.empty_like(self.rot_gt)
            self._st = torch.empty_like(self.trans_gt)
        return self._sq

    def _scratch_t(self):
        self._scratch_q()
        return self._st


LOOP_PACKED_KEYS = ("image_observed", "depth_gt_observed", "mask_gt_observed", "mask_observed")
LOOP_SMALL_KEYS = ("src_pose", "tgt_pose", "class_index", "mask_idx", "point_cloud_model",
                   "point_cloud_weights")


def pack_refine_batch(host, depth_factor=1000.0, pixel_means_bgr=None, pin=True,
                      mask_policy="upload"):
    """A RefineLoop batch (float32 blobs, make_refine_batch / the reference's DataBatch) in the
    dataset's STORAGE formats, as the files decode to before the reference's loaders touch them
    (lib/utils/image.py:186-305): colour uint8 [B,H,W,3] BGR, depth 16-bit in 1/depth_factor m
    (int16 view of the uint16 data), label image uint8 [B,H,W] + mask_idx, mask_observed uint8.
    The point cloud is the sampled MODEL cloud only; the observed cloud (R_gt X + t_gt,
    image.py:592-596) is computed on the device.  mask_policy "box_gt" (TRAIN.INIT_MASK of the
    shipped configs, image.py:330-343): mask_observed is the bounding rectangle of the label
    mask, which the reference's loader derives from the label file -- it is then not part of
    the packed batch at all and LoopFeed derives it on the device (rc_mask_box_dev).
    Returns CPU tensors (page-locked if `pin`)."""
    if mask_policy not in ("upload", "box_gt"):
        raise ValueError("mask_policy: 'upload' or 'box_gt'")
    means = np.asarray(synth.PIXEL_MEANS if pixel_means_bgr is None else pixel_means_bgr,
                       np.float64).reshape(-1)
    means_rgb = means[::-1].reshape(1, 3, 1, 1)
    rgb = np.clip(np.rint(host["image_observed"].astype(np.float64) + means_rgb), 0, 255).astype(np.uint8)
    out = {"image_observed": np.ascontiguousarray(rgb[:, ::-1].transpose(0, 2, 3, 1)),
           "depth_gt_observed": np.ascontiguousarray(
               np.rint(host["depth_gt_observed"][:, 0].astype(np.float64) * depth_factor)
               .astype(np.uint16).view(np.int16)),
           "mask_gt_observed": np.ascontiguousarray(host["mask_gt_observed"][:, 0].astype(np.uint8)),
           "mask_observed": np.ascontiguousarray(host["mask_observed"][:, 0].astype(np.uint8)),
           "mask_idx": np.ones((host["src_pose"].shape[0],), np.int32),
           "src_pose": np.ascontiguousarray(host["src_pose"], np.float32),
           "tgt_pose": np.ascontiguousarray(host["tgt_pose"], np.float32),
           "class_index": np.ascontiguousarray(host["class_index"], np.int32),
           "point_cloud_model": np.ascontiguousarray(host["point_cloud_model"], np.float32),
           "point_cloud_weights": np.ascontiguousarray(host["point_cloud_weights"], np.float32)}
    if mask_policy == "box_gt":
        del out["mask_observed"]
    res = {}
    for k, v in out.items():
        t = torch.from_numpy(v)
        res[k] = t.pin_memory() if pin else t
    return res


class LoopFeed(object):
    """End-to-end driver of a RefineLoop from HOST batches in the dataset's storage formats
    (pack_refine_batch): per batch

        H2D (copy stream, page-locked source; overlaps the previous batch's loop)
        -> on-device decode into the loop's persistent float32 blobs (prep.py: image.transform,
           DEPTH_FACTOR, label == idx, R_gt X + t_gt) -> iteration-0 labels + observed crop box
        -> `iterations` refinement iterations (one CUDA-graph replay)
        -> D2H of the refined poses and the last iteration's point-matching loss sums.

    The observed side is uploaded ONCE per loop and the rendered side never crosses PCIe (the
    device renders it) -- the reference re-uploads every float32 blob each iteration
    (batch_updater_py_multi.py:357-365, 12.4 MB per pair).  `depth` raw staging sets on the
    device; results of submit() number n are in pose_out[n % depth] / loss_out[n % depth]
    (page-locked) once done[n % depth] has completed."""

    def __init__(self, loop, example, iterations=4, depth_factor=1000.0, pixel_means_bgr=None,
                 depth=2, use_graph=True):
        """`example`: one packed batch (pack_refine_batch).  If it carries no "mask_observed"
        (mask_policy "box_gt") the observed mask is derived on the device from the label mask,
        as the reference's loader derives it on the host (lib/utils/image.py:330-343)."""
        from . import prep

        self.prep = prep
        self.loop, self.iterations, self.depth = loop, int(iterations), int(depth)
        self.depth_factor = float(depth_factor)
        self.means_bgr = np.asarray(
            synth.PIXEL_MEANS if pixel_means_bgr is None else pixel_means_bgr, np.float64).reshape(-1)
        dev, B = loop.device, loop.B
        self.box_gt = "mask_observed" not in example
        self.keys = tuple(k for k in LOOP_PACKED_KEYS + LOOP_SMALL_KEYS
                          if not (self.box_gt and k == "mask_observed"))
        if self.box_gt:
            self._box_ws = torch.zeros((B, _lib.RC_BOX_WS_INTS), dtype=torch.int32, device=dev)
            self._box_bbox = torch.empty((B, 4), dtype=torch.int32, device=dev)
        for k in self.keys:
            if example[k].shape[0] != B:
                raise ValueError("%s: batch %d, loop was built for %d" % (k, example[k].shape[0], B))
        self.raw = [{k: torch.empty_like(example[k], device=dev) for k in self.keys}
                    for _ in range(self.depth)]
        self.copy_stream = torch.cuda.Stream(device=dev)
        self.uploaded = [torch.cuda.Event() for _ in range(self.depth)]
        self.done = [torch.cuda.Event() for _ in range(self.depth)]
        self.pose_out = torch.empty((self.depth, B, 3, 4), dtype=torch.float32).pin_memory()
        self.loss_out = torch.empty((self.depth, B), dtype=torch.float32).pin_memory()
        self.n = 0
        self.use_graph = bool(use_graph)
        self.h2d_bytes = int(sum(example[k].numel() * example[k].element_size()
                                 for k in self.keys))
        self.d2h_bytes = int(B * 12 * 4 + B * 4)
        self.decode_launches = 8  # image, depth, 2 masks, points, labels, observed box, box copy

    def _decode(self, raw):
        p, lp = self.prep, self.loop
        o = lp.obs
        p.transform_batch(raw["image_observed"], self.means_bgr, out=o["image_observed"])
        p.decode_depth(raw["depth_gt_observed"], self.depth_factor, out=o["depth_gt_observed"])
        p.decode_mask(raw["mask_gt_observed"], raw["mask_idx"], out=o["mask_gt_observed"])
        if self.box_gt:
            mask_box(o["mask_gt_observed"], out=o["mask_observed"], bbox=self._box_bbox,
                     workspace=self._box_ws)
        else:
            p.decode_mask(raw["mask_observed"], raw["mask_idx"], out=o["mask_observed"])
        o["tgt_pose"].copy_(raw["tgt_pose"], non_blocking=True)
        o["point_cloud_model"].copy_(raw["point_cloud_model"], non_blocking=True)
        o["point_cloud_weights"].copy_(raw["point_cloud_weights"], non_blocking=True)
        p.get_point_cloud_observed(o["point_cloud_model"], o["tgt_pose"],
                                   out=o["point_cloud_observed"])
        lp.cls.copy_(raw["class_index"], non_blocking=True)
        lp.pose0.copy_(raw["src_pose"], non_blocking=True)
        lp.finish_batch()

    def submit(self, host):
        """Enqueue one batch (dict of page-locked CPU tensors from pack_refine_batch).  Returns
        the slot index; no host synchronisation."""
        i = self.n % self.depth
        lp = self.loop
        compute = torch.cuda.current_stream(lp.device)
        raw = self.raw[i]
        with torch.cuda.stream(self.copy_stream):
            if self.n >= self.depth:  # the slot's previous batch has been decoded and consumed
                self.copy_stream.wait_event(self.done[i])
            for k in self.keys:
                raw[k].copy_(host[k], non_blocking=True)
            self.uploaded[i].record(self.copy_stream)
        compute.wait_event(self.uploaded[i])
        self._decode(raw)
        if self.use_graph:
            if lp.graph is None or getattr(lp, "graph_iterations", None) != self.iterations:
                lp.capture(None, self.iterations)  # first batch: runs the loop eagerly + captures
            lp.replay()
        else:
            lp.run(self.iterations)
        self.pose_out[i].copy_(lp.pose[lp.cur], non_blocking=True)
        self.loss_out[i].copy_(lp.step.loss_sum, non_blocking=True)
        self.done[i].record(compute)
        self.n += 1
        return i


def make_refine_batch(meshes, batch, seed, K=synth.K_LINEMOD, num_points=synth.NUM_3D_SAMPLE,
                      renderer=None, occlusion=False, render_fn=None, mask_policy="hole"):
    """Synthetic observed side for RefineLoop: target renders come from `renderer`
    (a DeviceRenderer; the product's own rasteriser) so that no oracle is involved.
    `render_fn(cls [B] int32, tgt [B,3,4], pixel_means) -> (depth [B,1,H,W], image [B,3,H,W])`
    replaces the renderer where there is no GPU (the CPU arm of bench.py passes its own)."""
    rng = np.random.default_rng(seed)
    B = batch
    cls = rng.integers(0, len(meshes), size=B).astype(np.int32)
    src = np.zeros((B, 3, 4), np.float32)
    tgt = np.zeros((B, 3, 4), np.float32)
    pts = np.zeros((B, 3, num_points), np.float32)
    wts = np.zeros((B, 3, num_points), np.float32)
    obs = np.zeros((B, 3, num_points), np.float32)
    for b in range(B):
        src[b], tgt[b] = synth.random_pose_pair(rng, K)
        pts[b], wts[b] = synth.mesh_points(meshes[cls[b]], rng, num_points)
        obs[b] = tgt[b, :, :3].astype(np.float64) @ pts[b].astype(np.float64) + tgt[b, :, 3:].astype(np.float64)
    pm = synth.PIXEL_MEANS[[2, 1, 0]]
    if render_fn is not None:
        depth, image = render_fn(cls, tgt, pm)
        depth = np.ascontiguousarray(depth, np.float32)
        image = np.ascontiguousarray(image, np.float32)
    else:
        dev = renderer.device
        depth = np.empty((B, 1, synth.H, synth.W), np.float32)
        image = np.empty((B, 3, synth.H, synth.W), np.float32)
        chunk = max(1, min(B, renderer.max_batch))
        for lo in range(0, B, chunk):
            hi = min(B, lo + chunk)
            out = renderer.render_batch(torch.from_numpy(cls[lo:hi]).to(dev),
                                        torch.from_numpy(tgt[lo:hi]).to(dev),
                                        pixel_means=pm, want=("depth", "image"))
            torch.cuda.synchronize(dev)
            renderer.check_status()
            depth[lo:hi] = out["depth"].cpu().numpy()
            image[lo:hi] = out["image"].cpu().numpy()
    depth = (np.round(depth.astype(np.float64) * 1000.0) / 1000.0).astype(np.float32)  # 1 mm PNG
    # mask_gt_observed is the object's LABEL mask (label == cls_idx, lib/dataset/LM6D_REFINE.py),
    # so a nearer occluder removes pixels from it but never adds any; depth_gt_observed is the
    # sensor depth and does contain the occluder (Occlusion-LINEMOD-shaped)
    mask_gt = (depth > 0).astype(np.float32)
    if occlusion:
        for b in range(B):
            x0, y0 = rng.integers(0, synth.W - 80), rng.integers(0, synth.H - 80)
            depth[b, 0, y0:y0 + 80, x0:x0 + 80] = 0.4
            mask_gt[b, 0, y0:y0 + 80, x0:x0 + 80] = 0
    bg = rng.uniform(1, 255, size=image.shape).astype(np.float32) - pm.reshape(1, 3, 1, 1)
    image = np.where(mask_gt[:, :1] > 0, image, bg).astype(np.float32)
    # mask_policy "hole": the label mask with a 40x40 hole (a mask that differs from the label);
    # "box_gt": TRAIN.INIT_MASK of the shipped configs -- the label mask's bounding rectangle with
    # the reference's exclusive slice end (lib/utils/image.py:330-343)
    if mask_policy == "box_gt":
        mask_obs = np.zeros_like(mask_gt)
        for b in range(B):
            ys, xs = np.nonzero(mask_gt[b, 0])
            if len(ys):
                mask_obs[b, 0, ys.min():ys.max(), xs.min():xs.max()] = 1
    elif mask_policy == "hole":
        mask_obs = mask_gt.copy()
        for b in range(B):
            y0, x0 = rng.integers(0, synth.H - 40), rng.integers(0, synth.W - 40)
            mask_obs[b, 0, y0:y0 + 40, x0:x0 + 40] = 0
    else:
        raise ValueError("mask_policy: 'hole' or 'box_gt'")
    from . import pose as P

    return {"class_index": cls, "src_pose": src, "tgt_pose": tgt, "image_observed": image,
            "depth_gt_observed": depth, "mask_gt_observed": mask_gt, "mask_observed": mask_obs,
            "point_cloud_model": pts, "point_cloud_weights": wts, "point_cloud_observed": obs,
            "Kinv": P.make_Kinv(K)}
