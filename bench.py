#!/usr/bin/env python
"""bench.py -- render-compare pairs/s at 640x480 (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this framework
    python bench.py --impl reference --steps K --warmup W    # CPU reference arm (oracle port)

DEFAULT workload = BASELINE.json configs[4] ("cfg5"), the configuration the metric is quoted on
at 1/2/4/8 B200 and the one that exercises the WHOLE path north_star names:

    1024 pairs (ModelNet-unseen-shaped, 16 varied 20k-triangle meshes), strong-scaled over the
    N ranks (1024 / N pairs per GPU, contiguous blocks, no collective on the data path);
    per refinement iteration: network stand-in -> pose update (RT_transform, calc_RT_delta,
    calc_se3, KT) -> rasterise at the refined pose (+ fused image post / mask) -> GT flow ->
    crop box / zoom factor -> zoom-in of 13 planes -> point-matching loss forward + backward
    (+ invZoomTrans backward).  Reference: deepim/core/module.py:1205-1213 ->
    lib/pair_matching/batch_updater_py_multi.py:100-387.

A STEP is one refinement iteration over the rank's batch; a loop is 4 steps (TRAIN_ITER_SIZE).
A "pair" of the metric is one rendered/observed 640x480 pair through one iteration (SURVEY.md
8(d)).  --config 2|3|4 select the smaller BASELINE configs (2 = compare step only: flow + zoom +
loss, no rasteriser).

Printed JSON (one line, rank 0):
  value         pairs/s with every input resident in HBM: CUDA-graph replays of the loop, CUDA
                events on the launching stream, max over ranks; the K steps are repeated until
                >= 1 s of GPU time lies inside ONE event pair (`repeats`; `burst` = best block)
  e2e           the same through pipeline.LoopFeed from HOST buffers in the dataset's storage
                formats: H2D of the observed side (once per loop), on-device decode, 4 iterations,
                D2H of refined poses + loss sums -- all inside the timed region
  roofline      the dominant kernel (the bilinear resampler), timed by event-record nodes INSIDE
                the replayed graphs of the timed region; frac = SURVEY 8(d) algorithmic bytes,
                frac_compulsory = bytes the launch must move (outputs + crop-window reads)
  cpu_baseline  the oracle port of the same iteration (C rasteriser + C flow_kernel + numpy
                pose / zoom / loss) on the host cores, bounded sample; `legs` = BASELINE.md
                section 3's C1 / C4 / C5 / G0 single-stage baselines
"""
from __future__ import annotations

import argparse
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "render-compare pairs/s @640x480"
UNIT = "pairs/s"
BASE_SEED = 20181804  # default_rng(20181804 + cfg_id), + rank per GPU
ITERS = 4             # TRAIN_ITER_SIZE: refinement iterations per loop

CONFIGS = {
    2: dict(kind="step", batch=16, scaling="weak",
            name="cfg2: batch 16 LINEMOD-shaped pairs per GPU: GT flow + zoom-in (13 planes) + PM loss"),
    3: dict(kind="loop", batch=64, total=None, level=4, meshes=1, occlusion=False, backward=False,
            scaling="weak",
            name="cfg3: full 4-iteration refine loop (pose update, rasterise ~5k-tri mesh, flow, "
                 "zoom, loss), batch 64 per GPU"),
    4: dict(kind="loop", batch=None, total=256, level=4, meshes=8, occlusion=True, backward=False,
            scaling="strong",
            name="cfg4: Occlusion-LINEMOD-shaped full refine loop, 8 meshes, 256 pairs sharded "
                 "over the GPUs"),
    5: dict(kind="loop", batch=None, total=1024, level=5, meshes=16, occlusion=False, backward=True,
            scaling="strong",
            name="cfg5: ModelNet-unseen-shaped full refine loop (pose update, rasterise varied "
                 "20k-tri meshes, flow, zoom, PM loss fwd+bwd), 1024 pairs sharded over the GPUs"),
}
DEFAULT_CONFIG = 5
NSETS = 3  # cfg2: rotating input sets


def config_dict(config, n_gpus):
    """The `config` object of the JSON line: the WORKLOAD only, identical for both arms
    (--impl ours / reference) so that the driver compares like with like; what is specific to an
    arm's execution goes into the line's `run` object."""
    from mx_deepim_b200 import synth

    cfg = CONFIGS[config]
    d = {"workload": cfg["name"], "height": synth.H, "width": synth.W,
         "num_3d_sample": synth.NUM_3D_SAMPLE,
         "pair": "one rendered/observed 640x480 pair through ONE pass of the path "
                 "(one refinement iteration for the loop configs)",
         "parallelism": "dp%d (pairs sharded over the GPUs, no collective on the path)" % n_gpus,
         "l2": "inputs larger than L2: every step streams far more than the 126 MB L2 "
               "(cfg2: three rotating input sets, 651 MB per step; loop configs: >= 30 MB per pair "
               "per iteration)"}
    if cfg["kind"] == "loop":
        total = cfg["total"] if cfg["total"] is not None else cfg["batch"] * n_gpus
        d.update({"pairs_total": total, "batch_per_gpu": total // max(n_gpus, 1),
                  "iterations_per_loop": ITERS, "meshes": cfg["meshes"],
                  "triangles_per_mesh": 20 * 4 ** cfg["level"], "backward": cfg["backward"],
                  "occlusion": cfg["occlusion"],
                  "mask_observed": "TRAIN.INIT_MASK box_gt (bounding rectangle of the label mask)"})
    else:
        d.update({"pairs_total": cfg["batch"] * n_gpus, "batch_per_gpu": cfg["batch"]})
    return d


def _peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def _traffic(kernel, grid):
    """DRAM bytes per launch of `kernel` from the committed ncu --set full capture
    (profiles/dram_traffic.json, written by scripts/ncu_summary.py), if it was captured on the
    same launch shape; else None."""
    path = os.path.join(ROOT, "profiles", "dram_traffic.json")
    try:
        recs = json.load(open(path))
        for key in (kernel + "@%d" % int(grid), kernel):
            rec = recs.get(key)
            if rec and int(rec["grid"]) == int(grid):
                return rec["bytes_per_launch"], rec["source"]
    except Exception:
        pass
    return None, None


class ClockSampler(object):
    """nvidia-smi clocks / throttle reasons sampled every 100 ms while the GPU is under load."""

    FIELDS = ("clocks.sm,clocks.max.sm,utilization.gpu,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []
        self.windows = []  # (t0, t1) wall-clock windows of the timed regions

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.FIELDS,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smax, reasons, timed = [], [], set(), []
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for ts, line in self.lines:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                clk, mx, util = float(parts[0]), float(parts[1]), float(parts[2])
            except ValueError:
                continue
            smax.append(mx)
            in_window = any(a <= ts <= b + 0.15 for a, b in self.windows)
            if in_window:
                timed.append(clk)
            if util > 0 or in_window:
                sm.append(clk)
                for n, v in zip(names, parts[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
        return {"sm_mhz": statistics.median(timed) if timed else (statistics.median(sm) if sm else None),
                "sm_max_mhz": max(smax) if smax else None, "reasons": sorted(reasons),
                "samples_in_timed_regions": len(timed), "samples_under_load": len(sm),
                "samples": len(self.lines)}


# =====================================================================================
#  CPU reference arm (oracle port).  The ONLY code in this file that touches oracle/.
# =====================================================================================
_CPU = {}


def _oracle_render_fn(meshes):
    """render_fn for pipeline.make_refine_batch where there is no GPU: the C rasteriser oracle."""
    import numpy as np

    from mx_deepim_b200 import synth
    from oracle import raster_np

    def fn(cls, tgt, pm):
        B = len(cls)
        depth = np.zeros((B, 1, synth.H, synth.W), np.float32)
        image = np.zeros((B, 3, synth.H, synth.W), np.float32)
        for b in range(B):
            bgr, d = raster_np.render_py(meshes[int(cls[b])], tgt[b, :, :3], tgt[b, :, 3],
                                         synth.K_LINEMOD, r_type="mat", width=synth.W, height=synth.H,
                                         zNear=synth.ZNEAR, zFar=synth.ZFAR)
            image[b] = bgr[:, :, [2, 1, 0]].transpose(2, 0, 1) - np.asarray(pm, np.float32).reshape(3, 1, 1)
            depth[b, 0] = d
        return depth, image

    return fn


def _cpu_loop_pair(b):
    """The reference's CPU path for ONE pair through ONE refinement iteration (oracle port):
    stand-in prediction, pose update + labels + KT (numpy restatement of
    batch_updater_py_multi.py:211-312), render (C restatement of the GL pipeline), image post,
    flow_kernel (C), the zoom plugins (numpy + restated MXNet sampler), Transform3D + PM loss
    forward [+ backward + invZoomTrans backward]."""
    import numpy as np

    from mx_deepim_b200 import synth
    from oracle import pose_np, transform3d_np, updater_np, zoom_np

    host, meshes, backward = _CPU["host"], _CPU["meshes"], _CPU["backward"]
    s = slice(b, b + 1)
    zero, one = np.zeros(3, np.float32), np.ones(3, np.float32)
    src, tgt = host["src_pose"][s], host["tgt_pose"][s]
    q, t = pose_np.calc_RT_delta(src[0], tgt[0], zero, one, "CAMERA", "QUAT")
    rot_est = 0.5 * np.array([1, 0, 0, 0], np.float32) + 0.5 * np.asarray(q, np.float32)
    rot_est = (rot_est / np.linalg.norm(rot_est)).astype(np.float32)[None]
    trans_est = (0.5 * np.asarray(t, np.float32))[None]
    pm_bgr_order = synth.PIXEL_MEANS[[2, 1, 0]]
    ref = updater_np.updater_forward(meshes, host["class_index"][s], src, tgt, rot_est, trans_est,
                                     host["depth_gt_observed"][s], synth.K_LINEMOD, pm_bgr_order,
                                     zero, one, flow_threads=1)
    zm = zoom_np.zoom_mask_forward(host["mask_observed"][s], host["mask_gt_observed"][s],
                                   ref["depth_rendered"], ref["src_pose"], synth.K_LINEMOD)
    zi = zoom_np.zoom_image_with_factor_forward(zm[3], host["image_observed"][s],
                                                ref["image_rendered"], synth.PIXEL_MEANS[::-1])
    zf = zoom_np.zoom_flow_forward(zm[3], ref["flow"], ref["flow_weights"])
    zt = zoom_np.zoom_trans_forward(zm[3], ref["trans"])
    args = (host["point_cloud_model"][s], host["point_cloud_observed"][s],
            host["point_cloud_weights"][s], rot_est, trans_est, src, zero, one)
    _, _, sums = transform3d_np.pm_loss_forward(*args)
    chk = float(sums[0]) + float(zi[0][0, 0, 0, 0]) + float(zf[0][0, 0, 0, 0]) + float(zt[0, 0])
    if backward:
        g = transform3d_np.pm_loss_backward(*args, 1.0 / synth.NUM_3D_SAMPLE)
        gz = zoom_np.zoom_trans_backward(zm[3], np.asarray(g[1], np.float32).reshape(1, 3),
                                         b_inv_zoom=True, b_zoom_grad=False)
        chk += float(np.asarray(g[0]).reshape(-1)[0]) + float(gz[0, 0])
    return chk


def _cpu_step_pair(b):
    """cfg2: flow_kernel in C, numpy zoom plugins, numpy Transform3D + loss for ONE pair."""
    import numpy as np

    from mx_deepim_b200 import synth
    from oracle import flow_np, transform3d_np, zoom_np

    host = _CPU["host"]
    s = slice(b, b + 1)
    flow, valid = flow_np.flow_kernel_cpu(host["depth_rendered"][s], host["depth_gt_observed"][s],
                                          host["KT"][s], host["Kinv"], num_threads=1)
    zm = zoom_np.zoom_mask_forward(host["mask_observed"][s], host["mask_gt_observed"][s],
                                   host["depth_rendered"][s], host["src_pose"][s], synth.K_LINEMOD)
    zi = zoom_np.zoom_image_with_factor_forward(zm[3], host["image_observed"][s],
                                                host["image_rendered"][s], synth.PIXEL_MEANS[::-1])
    zf = zoom_np.zoom_flow_forward(zm[3], flow, np.tile(valid, [1, 2, 1, 1]))
    zt = zoom_np.zoom_trans_forward(zm[3], host["trans_gt"][s])
    _, _, sums = transform3d_np.pm_loss_forward(
        host["point_cloud_model"][s], host["point_cloud_observed"][s], host["point_cloud_weights"][s],
        host["rot_est"][s], host["trans_est"][s], host["src_pose"][s], np.zeros(3, np.float32),
        np.ones(3, np.float32))
    return float(sums[0]) + float(zi[0][0, 0, 0, 0]) + float(zf[0][0, 0, 0, 0]) + float(zt[0, 0])


class CpuReference(object):
    """Times the oracle port on the host cores: `pairs` pairs per step, one pair per process."""

    def __init__(self, config, host=None, meshes=None):
        import multiprocessing as mp

        import numpy as np

        from mx_deepim_b200 import pipeline, synth
        from oracle import build as obuild

        obuild.build_oracle()
        cfg = CONFIGS[config]
        self.cfg = cfg
        self.cores = os.cpu_count() or 1
        self.pairs = min(16, self.cores)
        if cfg["kind"] == "step":
            host = synth.make_compare_batch(self.pairs, BASE_SEED + config)
            self.fn = _cpu_step_pair
            self.what = "C flow_kernel + numpy zoom plugins + numpy Transform3D/loss"
        else:
            if meshes is None:
                rng = np.random.default_rng(BASE_SEED + config)
                meshes = [synth.make_mesh(rng, level=cfg["level"]) for _ in range(cfg["meshes"])]
            if host is None:
                host = pipeline.make_refine_batch(meshes, self.pairs, BASE_SEED + config,
                                                  occlusion=cfg["occlusion"], mask_policy="box_gt",
                                                  render_fn=_oracle_render_fn(meshes))
            else:
                self.pairs = min(self.pairs, host["src_pose"].shape[0])
            self.fn = _cpu_loop_pair
            self.what = ("numpy pose update/labels/KT + C rasteriser (GL restatement) + C flow_kernel "
                         "+ numpy zoom plugins + numpy Transform3D/loss%s"
                         % (" fwd+bwd" if cfg["backward"] else ""))
        _CPU.update(host=host, meshes=meshes, backward=bool(cfg.get("backward")))
        self.procs = min(self.cores, self.pairs)
        self.pool = mp.get_context("fork").Pool(self.procs)
        self.work = list(range(self.pairs))

    def step(self):
        t0 = time.perf_counter()
        self.pool.map(self.fn, self.work, chunksize=1)
        return time.perf_counter() - t0

    def sample(self):
        return ("%d pairs x 1 refinement iteration per step (1 pair per process, %d processes; host "
                "has %d cores); oracle port: %s" % (self.pairs, self.procs, self.cores, self.what))

    def close(self):
        self.pool.close()
        self.pool.join()


def baseline_legs(config, meshes, host, budget_s=3.0):
    """BASELINE.md section 3 single-stage baselines, each on a bounded sample (<= budget_s):
      C1  calc_flow numpy rule (lib/pair_matching/flow.py:12-81; the oracle restatement, pinned bit
          for bit to the reference's own calc_flow by tests/golden/flow_calc_flow_seed*.npz), 1 thread
      C4  torch-CPU grid_sample of the 13 planes, all cores (independent twin of the MXNet sampler)
      C5  CPU rasteriser (oracle/raster_oracle.c, the only possible CPU twin of
          render_py_multi.py:90-147), 1 thread per pair
    Returns {leg: {...}}.  /root/reference does not exist on the GPU box, hence the ports."""
    import numpy as np
    import torch

    from mx_deepim_b200 import synth
    from oracle import flow_np, raster_np

    legs = {}
    K = synth.K_LINEMOD

    def timed(fn, unit_count):
        fn()
        n, t0 = 0, time.perf_counter()
        while True:
            fn()
            n += 1
            dt = time.perf_counter() - t0
            if dt >= budget_s or n >= 50:
                return unit_count * n / dt, n

    try:
        src, tgt = host["src_pose"][0], host["tgt_pose"][0]
        d_obs = host["depth_gt_observed"][0, 0]
        v, n = timed(lambda: flow_np.calc_flow(d_obs, tgt, src, K, d_obs), 1)
        legs["C1_calc_flow_numpy"] = {"value": v, "unit": "pairs/s", "cores": 1, "reps": n,
                                      "what": "oracle.flow_np.calc_flow = restatement of "
                                              "lib/pair_matching/flow.py:12-81, one 640x480 pair"}
    except Exception as exc:  # pragma: no cover
        legs["C1_calc_flow_numpy"] = {"error": repr(exc)}
    try:
        x = torch.from_numpy(np.ascontiguousarray(
            np.concatenate([host["image_observed"][:1]] * 4 + [host["mask_observed"][:1]], axis=1)))
        theta = torch.tensor([[[0.35, 0.0, 0.05], [0.0, 0.35, -0.02]]], dtype=torch.float32)

        def gs():
            grid = torch.nn.functional.affine_grid(theta, list(x.shape), align_corners=True)
            return torch.nn.functional.grid_sample(x, grid, mode="bilinear", padding_mode="zeros",
                                                   align_corners=True)

        v, n = timed(gs, 1)
        legs["C4_torch_cpu_grid_sample"] = {"value": v, "unit": "pairs/s", "reps": n,
                                            "cores": torch.get_num_threads(),
                                            "what": "torch.nn.functional.grid_sample, 13 planes of one pair"}
    except Exception as exc:  # pragma: no cover
        legs["C4_torch_cpu_grid_sample"] = {"error": repr(exc)}
    try:
        if meshes is not None:
            m = meshes[int(host["class_index"][0])]
            tp = host["tgt_pose"][0]
            v, n = timed(lambda: raster_np.render_py(m, tp[:, :3], tp[:, 3], K, r_type="mat"), 1)
            legs["C5_cpu_rasteriser"] = {"value": v, "unit": "pairs/s", "cores": 1, "reps": n,
                                         "triangles_per_s": v * len(m["triangles"]),
                                         "what": "oracle/raster_oracle.c, %d-triangle mesh, one pair"
                                                 % len(m["triangles"])}
    except Exception as exc:  # pragma: no cover
        legs["C5_cpu_rasteriser"] = {"error": repr(exc)}
    return legs


def g0_reference_flow_kernel(dev, B=16):
    """G0 kernel-only: the reference's UNMODIFIED flow_kernel (gpu_flow_kernel.cu:32-69, its own
    launch configuration, oracle/_ref built by oracle/ref_flow_shim.cu) against flow_kernel_v4 on
    the same device buffers, CUDA events, L2 flushed between launches.  A checker-side leg: it
    executes oracle/_ref, never the other way round."""
    import numpy as np
    import torch

    from mx_deepim_b200 import _lib, flow as F, synth
    from oracle import flow_np

    if not flow_np.reference_kernel_available():
        return {"unavailable": "oracle/_ref/libgpu_flow_ref.so not built"}
    try:
        host = synth.make_compare_batch(B, BASE_SEED + 99)
        t = lambda k: torch.from_numpy(np.ascontiguousarray(host[k])).to(dev)
        ds, dt, KT, Ki = t("depth_rendered"), t("depth_gt_observed"), t("KT"), t("Kinv")
        f_ref, v_ref = torch.zeros((B, 2, synth.H, synth.W), device=dev), torch.zeros((B, 1, synth.H, synth.W), device=dev)
        f_our, v_our = torch.empty_like(f_ref), torch.empty_like(v_ref)
        flush = torch.empty((256 << 20) // 4, dtype=torch.float32, device=dev)
        sp = _lib.stream_ptr(None)

        def ref():
            flow_np.reference_kernel_dev((ds.data_ptr(), tuple(ds.shape)), dt.data_ptr(), KT.data_ptr(),
                                         Ki.data_ptr(), f_ref.data_ptr(), v_ref.data_ptr(), sp)

        def ours():
            F.gpu_flow_dev(ds, dt, KT, Ki, flow=f_our, valid=v_our)

        out = {}
        for name, fn in (("reference_flow_kernel", ref), ("flow_kernel_v4", ours)):
            fn()
            torch.cuda.synchronize(dev)
            ts = []
            for _ in range(10):
                flush.fill_(1.0)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                fn()
                e1.record()
                torch.cuda.synchronize(dev)
                ts.append(e0.elapsed_time(e1) * 1e3)
            us = statistics.median(ts)
            out[name] = {"us": us, "gbs": 20.0 * synth.H * synth.W * B / (us * 1e-6) / 1e9}
        out["bit_identical"] = bool(torch.equal(f_ref, f_our) and torch.equal(v_ref, v_our))
        out["batch"] = B
        out["what"] = ("kernel-only, device buffers, median of 10 launches with a 256 MB L2 flush between; "
                       "reference = gpu_flow_kernel.cu:32-69 unmodified, grid CAFFE_GET_BLOCKS(count) x 512")
        return out
    except Exception as exc:  # pragma: no cover
        return {"error": repr(exc)}


def run_reference(args, rank):
    if rank != 0:
        return
    ref = CpuReference(args.config)
    for _ in range(max(min(args.warmup, 2), 1)):
        ref.step()
    times = [ref.step() for _ in range(args.steps)]
    ref.close()
    total = sum(times)
    value = ref.pairs * args.steps / total
    cfg = CONFIGS[args.config]
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True,
        "scaling": cfg["scaling"], "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_dict(args.config, args.gpus),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": ref.procs, "kind": "port",
                         "sample": ref.sample()},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# =====================================================================================
#  GPU arm, refine-loop configs (3, 4, 5; 5 = default)
# =====================================================================================
def run_loop(args, rank, world, local_rank):
    import numpy as np
    import torch
    import torch.distributed as dist

    from mx_deepim_b200 import _lib, pipeline, shard, synth

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU fallback")
    cfg = CONFIGS[args.config]
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    numa_cores = shard.bind_to_gpu_numa(local_rank) if world > 1 else None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()

    if args.pairs_total:  # experiment knob (e.g. one GPU's share of an 8-GPU run); not the metric
        cfg = dict(cfg, total=args.pairs_total, name=cfg["name"] + " [--pairs-total %d]" % args.pairs_total)
    if cfg["total"] is not None:
        lo, hi = shard.partition(cfg["total"], world, rank)
        B = hi - lo
        if cfg["total"] % world:
            raise SystemExit("bench.py: %d pairs do not split evenly over %d GPUs" % (cfg["total"], world))
    else:
        B = cfg["batch"]
    rng = np.random.default_rng(BASE_SEED + args.config)
    meshes = [synth.make_mesh(rng, level=cfg["level"]) for _ in range(cfg["meshes"])]
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    loop = pipeline.RefineLoop(meshes, B, dev, seed=rank, backward=cfg["backward"])
    host = pipeline.make_refine_batch(meshes, B, shard.rank_seed(BASE_SEED + args.config, rank),
                                      renderer=loop.renderer, occlusion=cfg["occlusion"],
                                      mask_policy="box_gt")
    loop.set_batch(host)
    K = max(ITERS, ((args.steps + ITERS - 1) // ITERS) * ITERS)  # whole loops
    n_loops = K // ITERS
    Wm = max(args.warmup, 3)
    w_loops = (Wm + ITERS - 1) // ITERS

    # resampler marks (2 = before, 3 = after) as event-record nodes inside the captured loop:
    # they do not serialise the step (only the flow/bbox/loss marks do)
    marks = [[None, None, _lib.Event(), _lib.Event(), None, None] for _ in range(ITERS)]
    if args.no_graph:
        def run_loop_once():
            loop.set_batch_poses()
            loop.run(ITERS, marks=marks)
        issue = "eager launches (--no-graph)"
    else:
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            loop.capture(None, ITERS, marks=marks)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize()
        run_loop_once = loop.replay
        issue = "one CUDA graph per %d-iteration loop (RefineLoop.capture / replay)" % ITERS
    for _ in range(max(w_loops, 1)):
        run_loop_once()
    torch.cuda.synchronize()
    loop.check_status()

    # ------------------------------------------------ device-resident arm (`value`)
    e0, e1 = _lib.Event(), _lib.Event()

    def block(reps):
        barrier()
        torch.cuda.synchronize()
        t0 = time.time()
        e0.record()
        for _ in range(reps * n_loops):
            run_loop_once()
        e1.record()
        torch.cuda.synchronize()
        t1 = time.time()
        barrier()
        own = e0.elapsed_ms(e1)
        per_rank[:] = shard.gather_scalar(own, dev)
        return shard.reduce_max(own, dev), (t0, t1)

    per_rank = []                               # each rank's own device time of the last block
    ms_block, _ = block(1)                      # one K-step block: the burst figure
    repeats = max(1, int(math.ceil(args.min_seconds * 1e3 / max(ms_block, 1e-3))))
    if world > 1:  # every rank must agree on the repeat count
        t = torch.tensor([repeats], dtype=torch.int64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        repeats = int(t.item())
    ms_total, window = block(repeats)           # the sustained figure: >= min_seconds in one event pair
    sampler.windows.append(window)
    steps_timed = K * repeats
    ms_per_step = ms_total / steps_timed
    value = world * B / (ms_per_step * 1e-3)
    burst = world * B * K / (ms_block * 1e-3)
    # the resampler inside the LAST loop of the timed region (event nodes of the replayed graph)
    k_us = [m[2].elapsed_ms(m[3]) * 1e3 for m in marks]
    kernel_us = sum(k_us) / len(k_us)
    loop.check_status()
    loss_checksum = float(loop.step.loss_sum.double().sum().item())

    # ------------------------------------------------ end-to-end arm (`e2e`): host buffers
    packed = pipeline.pack_refine_batch(host, mask_policy="box_gt")
    packed2 = {k: v.flip(0).contiguous().pin_memory() for k, v in packed.items()}  # a second host batch
    feed = pipeline.LoopFeed(loop, packed, iterations=ITERS, use_graph=not args.no_graph)
    sets = [packed, packed2]
    for i in range(2):
        feed.submit(sets[i % 2])
    torch.cuda.synchronize()
    Le = max(2, min(n_loops * repeats, int(math.ceil(args.min_seconds * 1e3 / max(ms_block / n_loops, 1e-3) / 2))))
    if world > 1:
        t = torch.tensor([Le], dtype=torch.int64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        Le = int(t.item())
    barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    w0 = time.time()
    e0.record()
    for i in range(Le):
        feed.submit(sets[i % 2])
    e1.record()
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    sampler.windows.append((w0, time.time()))
    barrier()
    ms_e = shard.reduce_max(max(e0.elapsed_ms(e1), wall * 1e3), dev)  # wall clock includes the final sync
    e2e_value = world * B * ITERS * Le / (ms_e * 1e-3)
    pose_checksum = float(feed.pose_out[(feed.n - 1) % feed.depth].double().sum().item())
    loop.check_status()
    h2d_gbs = world * feed.h2d_bytes * Le / (ms_e * 1e-3) / 1e9

    # -------------------------------- per-kernel breakdown: separate SERIALISED pass (marks
    # between the launches turn the overlapping launches into normal ones), not timed
    n_mark = 2
    rm = [[[_lib.Event() for _ in range(_lib.RC_RENDER_MARKS)] for _ in range(ITERS)] for _ in range(n_mark)]
    sm = [[[_lib.Event() for _ in range(_lib.RC_STEP_MARKS)] for _ in range(ITERS)] for _ in range(n_mark)]
    for l in range(n_mark):
        loop.set_batch_poses()
        for it in range(ITERS):
            loop.iterate(it, marks=sm[l][it], render_marks=rm[l][it])
    torch.cuda.synchronize()
    flat_r = [m for l in rm for m in l]
    flat_s = [m for l in sm for m in l]
    r_names = ["raster_clear_vertex_setup", "raster_scan_clip", "raster_fill", "raster_cover_resolve"]
    kern = {n: sum(m[j].elapsed_ms(m[j + 1]) for m in flat_r) / len(flat_r) * 1e3
            for j, n in enumerate(r_names)}
    s_names = {"frame_tiles_flow": (0, 1), "zoom_factor_boxes": (1, 2), "zoom_sample": (2, 3),
               "pm_loss_fwd_bwd_zoomtrans": (4, 5)}
    for n, (a, b) in s_names.items():
        kern[n] = sum(m[a].elapsed_ms(m[b]) for m in flat_s) / len(flat_s) * 1e3
    stats = loop.renderer.stats(B)
    clocks = sampler.stop() if rank == 0 else None

    # ------------------------------------------------ CPU baseline (rank 0, N = 1 only)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        P = min(16, os.cpu_count() or 1, B)
        sub = {k: (v[:P] if isinstance(v, np.ndarray) and v.ndim > 0 and v.shape[0] == B else v)
               for k, v in host.items()}
        ref = CpuReference(args.config, host=sub, meshes=meshes)
        ref.step()
        reps, total = 0, 0.0
        while total < 10.0 and reps < 20:
            total += ref.step()
            reps += 1
        ref.close()
        cpu = {"value": ref.pairs * reps / total, "unit": UNIT, "cores": ref.procs, "kind": "port",
               "sample": "%d steps of %s" % (reps, ref.sample()),
               "legs": baseline_legs(args.config, meshes, sub)}
        cpu["legs"]["G0_reference_flow_kernel_only"] = g0_reference_flow_kernel(dev)

    if rank == 0:
        peak, peak_src = _peaks()
        px = synth.H * synth.W
        kbytes = pipeline.BYTES_PER_PX_ZOOM_SAMPLE * px * B  # every plane once
        achieved = kbytes / (kernel_us * 1e-6) / 1e9
        # what the launch MUST move: 13 output planes written + the crop windows read
        zf = loop.step.zoom_factor.cpu().numpy()
        win = np.clip(zf[:, 0], 0, 1) * np.clip(zf[:, 1], 0, 1)  # crop area / frame area
        comp = float(52 * px * B + 48 * px * float(win.sum()))   # 12 distinct source planes
        grid = ((synth.W + 127) // 128) * ((synth.H + 15) // 16) * B  # zoom_sample2_kernel: 128 x 16 tiles
        traffic, traffic_src = _traffic("zoom_sample2_kernel", grid)
        render_s = sum(kern[n] for n in r_names) * 1e-6
        ntri = sum(loop.renderer.n_triangles[c] for c in host["class_index"])
        fg = float((loop.depth_rendered > 0).sum().item())
        iter_bytes = (pipeline.algorithmic_bytes_per_pair(synth.H, synth.W, synth.NUM_3D_SAMPLE) + 20 * px) * B
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K,
            "warmup": Wm, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": cfg["scaling"], "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "repeats": repeats, "timed_ms": ms_total,
            "per_rank_ms_per_step": [t / steps_timed for t in per_rank],
            "burst": {"value": burst, "ms_per_step": ms_block / K, "steps": K},
            "config": config_dict(args.config, world),
            "run": {"issue": issue, "batch_per_gpu": B,
                    "triangles_per_mesh": len(meshes[0]["triangles"]),
                    "l2": "per-iteration algorithmic traffic %.0f MB per GPU > 126 MB L2" % (iter_bytes / 1e6)},
            "e2e": {"value": e2e_value, "unit": UNIT,
                    "h2d_bytes_per_step": feed.h2d_bytes // ITERS, "d2h_bytes_per_step": feed.d2h_bytes // ITERS,
                    "h2d_bytes_per_loop": feed.h2d_bytes, "d2h_bytes_per_loop": feed.d2h_bytes,
                    "loops": Le, "steps": Le * ITERS, "pose_checksum": pose_checksum,
                    "h2d_gbs_all_ranks": h2d_gbs,
                    "feed": "pipeline.LoopFeed: observed side in storage formats (u8 BGR image, u16 mm "
                            "depth, u8 label, model points, poses; mask_observed = TRAIN.INIT_MASK box_gt "
                            "derived from the label on the device, as the reference's loader derives it on "
                            "the host) from page-locked host memory, "
                            "uploaded ONCE per 4-iteration loop on a copy stream (overlaps the previous "
                            "loop), decoded on device; the rendered side never crosses PCIe; D2H of "
                            "refined poses + loss sums per loop",
                    "limiter": "host->device copy: %.1f GB/s aggregate over %d rank(s)" % (h2d_gbs, world),
                    "host_cores_bound": (len(numa_cores) if numa_cores else None)},
            "gpu_launches": loop.launches_per_iteration * steps_timed,
            "roofline": {"bound": "hbm", "kernel": "zoom_sample2_kernel", "achieved": achieved,
                         "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                         "kernel_us": kernel_us,
                         "kernel_us_source": "event-record nodes around the resampler inside the replayed "
                                             "graph: the %d launches of the last loop of the timed region" % ITERS,
                         "kernel_share_of_step": kernel_us * 1e-3 / ms_per_step,
                         "algorithmic_bytes_per_launch": kbytes,
                         "compulsory_bytes_per_launch": comp,
                         "frac_compulsory": comp / (kernel_us * 1e-6) / 1e9 / peak,
                         "dram_frac": (traffic / (kernel_us * 1e-6) / 1e9 / peak) if traffic else None,
                         "note": "frac counts every source plane in full (SURVEY 8(d)); a zoom-in reads "
                                 "only its crop window, so frac_compulsory (13 output planes + window "
                                 "reads) is the physical figure",
                         "step_achieved_gbs": iter_bytes / (ms_per_step * 1e-3) / 1e9},
            "rasteriser": {"serialised_us": render_s * 1e6,
                           "triangles_per_s": ntri / render_s if render_s > 0 else None,
                           "covered_px_per_s": fg / render_s if render_s > 0 else None,
                           "algorithmic_gbs": (20 * px * B) / render_s / 1e9 if render_s > 0 else None,
                           "records": stats["records"], "tile_entries": stats["tile_entries"],
                           "micro_direct": stats["direct"],
                           "covered_px_per_pair": fg / B},
            "kernel_us_serialised": kern,
            "loss_checksum": loss_checksum,
            "clocks": clocks,
            "cpu_baseline": cpu,
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


# =====================================================================================
#  GPU arm, cfg2: the compare step alone (flow + zoom + loss)
# =====================================================================================
def run_step(args, rank, world, local_rank):
    import numpy as np
    import torch
    import torch.distributed as dist

    from mx_deepim_b200 import _lib, pipeline, shard, synth

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU fallback")
    cfg = CONFIGS[2]
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    numa_cores = shard.bind_to_gpu_numa(local_rank) if world > 1 else None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()

    B, K, Wm = cfg["batch"], args.steps, max(args.warmup, 3)
    SEED = BASE_SEED + 2
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    hosts = [synth.make_compare_batch(B, SEED + 1000 * rank + s) for s in range(NSETS)]
    step = pipeline.CompareStep(B, dev)
    dsets = [step.to_device(h, non_blocking=False) for h in hosts]
    bound = [step.bind(x) for x in dsets]
    torch.cuda.synchronize()

    for i in range(Wm):
        step.run(None, args=bound[i % NSETS])
    torch.cuda.synchronize()
    marks = [[None, None, _lib.Event(), _lib.Event(), None, None] for _ in range(K)]
    e0, e1 = _lib.Event(), _lib.Event()

    def block(reps, use_marks):
        barrier()
        torch.cuda.synchronize()
        t0 = time.time()
        e0.record()
        for r in range(reps):
            for i in range(K):
                step.run(None, args=bound[i % NSETS], marks=marks[i] if use_marks else None)
        e1.record()
        torch.cuda.synchronize()
        t1 = time.time()
        barrier()
        return shard.reduce_max(e0.elapsed_ms(e1), dev), (t0, t1)

    ms_block, _ = block(1, True)
    repeats = max(1, int(math.ceil(args.min_seconds * 1e3 / max(ms_block, 1e-3))))
    if world > 1:
        t = torch.tensor([repeats], dtype=torch.int64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        repeats = int(t.item())
    ms_total, window = block(repeats, True)
    sampler.windows.append(window)
    ms_per_step = ms_total / (K * repeats)
    value = world * B / (ms_per_step * 1e-3)
    k_ms = [m[2].elapsed_ms(m[3]) for m in marks]  # the last repeat's K launches
    kernel_ms = sum(k_ms) / len(k_ms)

    full = [[_lib.Event() for _ in range(_lib.RC_STEP_MARKS)] for _ in range(min(K, 20))]
    for i, m in enumerate(full):
        step.run(None, args=bound[i % NSETS], marks=m)
    torch.cuda.synchronize()
    names = {"flow": (0, 1), "zoom_bbox_factor_trans": (1, 2), "zoom_sample": (2, 3), "pm_loss": (4, 5)}
    breakdown = {n: sum(m[a].elapsed_ms(m[b]) for m in full) / len(full) * 1e3
                 for n, (a, b) in names.items()}

    # end to end from the dataset's storage formats (uint8 BGR images, uint16 mm depth, uint8
    # masks in page-locked host memory -> H2D -> on-device decode -> step -> D2H of loss + zoom
    # factors); e2e_f32 = the same from float32 blobs as the reference's DataBatch holds them
    means_bgr = np.asarray(synth.PIXEL_MEANS, np.float64).reshape(-1)
    means_rgb = means_bgr[::-1].reshape(1, 3, 1, 1)

    def pack(h):
        out = {}
        for k in ("image_observed", "image_rendered"):
            rgb = np.clip(np.rint(h[k].astype(np.float64) + means_rgb), 0, 255).astype(np.uint8)
            out[k] = np.ascontiguousarray(rgb[:, ::-1].transpose(0, 2, 3, 1))  # BGR, HWC
        for k in ("depth_gt_observed", "depth_rendered"):
            out[k] = np.ascontiguousarray(np.rint(h[k][:, 0].astype(np.float64) * 1000.0)
                                          .astype(np.uint16).view(np.int16))
        for k in ("mask_observed", "mask_gt_observed"):
            out[k] = np.ascontiguousarray(h[k][:, 0].astype(np.uint8))
        return out

    pinned = [{k: torch.from_numpy(np.ascontiguousarray(h[k])).pin_memory()
               for k in pipeline.INPUT_KEYS} for h in hosts]
    packed = []
    for h, pf in zip(hosts, pinned):
        d = dict(pf)
        d.update({k: torch.from_numpy(v).pin_memory() for k, v in pack(h).items()})
        packed.append(d)

    def e2e_run(feed, sets):
        Ke = max(3, min(K * repeats, 400))
        for i in range(3):
            feed.submit(sets[i % NSETS])
        torch.cuda.synchronize()
        barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        w0 = time.time()
        e0.record()
        for i in range(Ke):
            feed.submit(sets[i % NSETS])
        e1.record()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        sampler.windows.append((w0, time.time()))
        barrier()
        ms = shard.reduce_max(max(e0.elapsed_ms(e1), wall * 1e3), dev)
        return world * B * Ke / (ms * 1e-3), Ke

    pfeed = pipeline.HostFeed(step, packed[0], packed=True, depth_factor=1000.0, pixel_means_bgr=means_bgr)
    e2e_value, Ke = e2e_run(pfeed, packed)
    ffeed = pipeline.HostFeed(step, pinned[0])
    e2e_f32, _ = e2e_run(ffeed, pinned)
    loss_host = float(pfeed.loss[(pfeed.n - 1) % pfeed.depth].sum())
    clocks = sampler.stop() if rank == 0 else None

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        ref = CpuReference(2)
        ref.step()
        reps, total = 0, 0.0
        while total < 10.0 and reps < 20:
            total += ref.step()
            reps += 1
        ref.close()
        cpu = {"value": ref.pairs * reps / total, "unit": UNIT, "cores": ref.procs, "kind": "port",
               "sample": "%d steps of %s" % (reps, ref.sample())}

    if rank == 0:
        peak, peak_src = _peaks()
        px = synth.H * synth.W
        kbytes = pipeline.BYTES_PER_PX_ZOOM_SAMPLE * px * B
        achieved = kbytes / (kernel_ms * 1e-3) / 1e9
        zf = step.zoom_factor.cpu().numpy()
        win = np.clip(zf[:, 0], 0, 1) * np.clip(zf[:, 1], 0, 1)
        comp = float(52 * px * B + 48 * px * float(win.sum()))
        grid = ((synth.W + 127) // 128) * ((synth.H + 15) // 16) * B  # zoom_sample2_kernel: 128 x 16 tiles
        traffic, traffic_src = _traffic("zoom_sample2_kernel", grid)
        step_bytes = pipeline.algorithmic_bytes_per_pair(synth.H, synth.W, synth.NUM_3D_SAMPLE) * B
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K,
            "warmup": Wm, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "repeats": repeats, "timed_ms": ms_total,
            "burst": {"value": world * B * K / (ms_block * 1e-3), "ms_per_step": ms_block / K, "steps": K},
            "config": config_dict(2, world),
            "run": {"l2": "%d rotating input sets; %.0f MB touched per step > 126 MB L2"
                          % (NSETS, step_bytes / 1e6)},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": pfeed.h2d_bytes,
                    "d2h_bytes_per_step": pfeed.d2h_bytes, "steps": Ke, "loss_checksum": loss_host,
                    "feed": "pipeline.HostFeed(packed): uint8 BGR HWC images, uint16 mm depth, uint8 masks "
                            "from page-locked host memory, decoded on device; upload of batch i+1 "
                            "overlapped with the step of batch i",
                    "host_cores_bound": (len(numa_cores) if numa_cores else None)},
            "e2e_f32": {"value": e2e_f32, "unit": UNIT, "h2d_bytes_per_step": ffeed.h2d_bytes,
                        "d2h_bytes_per_step": ffeed.d2h_bytes,
                        "feed": "float32 blobs as the reference's DataBatch holds them (12.4 MB per pair)"},
            "gpu_launches": step.launches_per_step * K * repeats,
            "roofline": {"bound": "hbm", "kernel": "zoom_sample2_kernel", "achieved": achieved,
                         "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                         "kernel_us": kernel_ms * 1e3,
                         "algorithmic_bytes_per_launch": kbytes,
                         "compulsory_bytes_per_launch": comp,
                         "frac_compulsory": comp / (kernel_ms * 1e-3) / 1e9 / peak,
                         "dram_frac": (traffic / (kernel_ms * 1e-3) / 1e9 / peak) if traffic else None,
                         "note": "frac counts every source plane in full (SURVEY 8(d)); a zoom-in reads "
                                 "only its crop window: frac_compulsory (13 output planes + window reads) "
                                 "is the physical figure",
                         "step_achieved_gbs": step_bytes / (ms_per_step * 1e-3) / 1e9},
            "kernel_us": breakdown,
            "clocks": clocks,
            "cpu_baseline": cpu,
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true",
                    help="refine-loop configs: issue every launch from the host instead of "
                         "replaying one CUDA graph per loop")
    ap.add_argument("--min-seconds", type=float, default=1.0,
                    help="repeat the K-step block until this much GPU time is inside one event pair")
    ap.add_argument("--pairs-total", type=int, default=0,
                    help="experiments only: override the config's total number of pairs")
    ap.add_argument("--config", type=int, default=DEFAULT_CONFIG, choices=sorted(CONFIGS),
                    help="BASELINE.json config (5 = default: the full refine loop the metric is "
                         "quoted on; 2 = compare step only)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        if args.steps > 30:
            args.steps = 30  # bounded: each step is ~0.5-1 s of CPU work per core
        run_reference(args, rank)
        return
    if CONFIGS[args.config]["kind"] == "step":
        run_step(args, rank, world, local_rank)
    else:
        run_loop(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
