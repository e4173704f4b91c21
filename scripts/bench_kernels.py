#!/usr/bin/env python
"""Per-kernel roofline table: every kernel of the path and of its neighbours (SURVEY.md 8(a) /
8(f)) timed alone with CUDA events on the launching stream, L2 flushed between repetitions,
achieved GB/s = ALGORITHMIC bytes / median time against the measured HBM peak.  Also times the
reference's own gpu_flow_kernel.cu (oracle/_ref, built unmodified for sm_100a) through its
host-pointer `_flow` next to ours.

    python scripts/bench_kernels.py [--batch 16] [--reps 30] > gpurun_out/kernels.json

Not the bench.py contract: this is the supporting table copied into profiles/.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np
import torch

from mx_deepim_b200 import _lib, evaluation, flow as flow_mod, pipeline, prep, synth
from mx_deepim_b200.operator_py import _common as C


def peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"])
    return 6650.0


def timed(fn, reps, flush):
    """median microseconds of fn() (one or more launches on the current stream)."""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    out = []
    for _ in range(reps):
        flush.add_(1.0)  # 512 MB read+write: evicts the 126 MB L2
        e0, e1 = _lib.Event(), _lib.Event()
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        out.append(e0.elapsed_ms(e1) * 1e3)
    return statistics.median(out)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--reps", type=int, default=30)
    a = ap.parse_args()
    B, reps = a.batch, a.reps
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    H, W, N = synth.H, synth.W, synth.NUM_3D_SAMPLE
    px = H * W
    flush = torch.zeros((64 << 20,), dtype=torch.float32, device=dev)
    host = synth.make_compare_batch(B, 20181804 + 2)
    step = pipeline.CompareStep(B, dev)
    x = step.to_device(host, non_blocking=False)
    args = step.bind(x)
    step.run(None, args=args)
    torch.cuda.synchronize()
    pk = peak()
    rows = []

    def row(name, us, nbytes, note=""):
        gbs = nbytes / (us * 1e-6) / 1e9
        rows.append({"kernel": name, "us": round(us, 2), "algorithmic_MB": round(nbytes / 1e6, 3),
                     "GBps": round(gbs, 1), "frac_of_hbm_peak": round(gbs / pk, 4), "note": note})

    lib = _lib.load()
    sp = _lib.stream_ptr
    # ---- the four kernels of the compare step
    row("flow_kernel_v4 (rc_flow_dev)", timed(lambda: flow_mod.gpu_flow_dev(
        x["depth_rendered"], x["depth_gt_observed"], x["KT"], x["Kinv"], step.flow, step.valid),
        reps, flush), 20 * px * B, "8 B/px read + 12 B/px written")
    row("zoom_bbox_kernel (rc_zoom_factor_dev)", timed(lambda: C.zoom_factor(
        x["mask_gt_observed"], x["depth_rendered"], x["src_pose"], step.K, None, step.zoom_factor,
        _lib.RC_ZF_MASK), reps, flush), 8 * px * B, "two planes read once")

    # the resampler alone: time it through the step's marks (serial mode gives clean marks)
    def step_marks():
        m = [_lib.Event() for _ in range(_lib.RC_STEP_MARKS)]
        step.run(None, args=args, marks=m)
        return m

    for _ in range(3):
        step_marks()
    ts = []
    for _ in range(reps):
        flush.add_(1.0)
        m = step_marks()
        torch.cuda.synchronize()
        ts.append(m[2].elapsed_ms(m[3]) * 1e3)
    row("zoom_sample_kernel (13 planes)", statistics.median(ts), 104 * px * B,
        "52 B/px read + 52 B/px written (algorithmic; DRAM reads are the crop window only)")
    row("pm_fwd_kernel (rc_pm_loss_fwd_dev)", timed(lambda: lib.rc_pm_loss_fwd_dev(
        x["point_cloud_model"].data_ptr(), x["point_cloud_observed"].data_ptr(),
        x["point_cloud_weights"].data_ptr(), x["rot_est"].data_ptr(), x["trans_est"].data_ptr(),
        x["src_pose"].data_ptr(), step.T_means, step.T_stds, step.rot_coord, step.loss_type,
        step.norm_term, 1.0, None, step.loss_map.data_ptr(), step.loss_sum.data_ptr(), B, N,
        sp()), reps, flush), 4 * 3 * N * 4 * B, "latency-bound: 0.14 MB per pair")
    t_step = timed(lambda: step.run(None, args=args), reps, flush)
    row("whole compare step (4 launches, overlapped)", t_step,
        pipeline.algorithmic_bytes_per_pair(H, W, N) * B)

    # ---- neighbours of the path (8(f))
    mb_out = torch.empty_like(x["mask_observed"])
    mb_box = torch.empty((B, 4), dtype=torch.int32, device=dev)
    mb_ws = torch.zeros((B, _lib.RC_BOX_WS_INTS), dtype=torch.int32, device=dev)
    row("mask_bbox + mask_box_fill (rc_mask_box_dev)", timed(lambda: pipeline.mask_box(
        x["mask_gt_observed"], out=mb_out, bbox=mb_box, workspace=mb_ws), reps, flush),
        8 * px * B, "one plane read, one written; 2 launches")
    im8 = torch.randint(0, 256, (B, H, W, 3), dtype=torch.uint8, device=dev)
    im_out = torch.empty((B, 3, H, W), dtype=torch.float32, device=dev)
    row("prep_image_kernel", timed(lambda: prep.transform_batch(im8, synth.PIXEL_MEANS, out=im_out),
                                   reps, flush), 15 * px * B, "3 B/px read + 12 B/px written")
    d16 = torch.randint(0, 3000, (B, H, W), dtype=torch.int16, device=dev)
    d_out = torch.empty((B, 1, H, W), dtype=torch.float32, device=dev)
    row("prep_depth_kernel", timed(lambda: prep.decode_depth(d16, 1000.0, out=d_out), reps, flush),
        6 * px * B, "2 B/px read + 4 B/px written")
    T = torch.from_numpy(np.stack([prep.make_transform(synth.K_LINEMOD, host["src_pose"][b],
                                                       host["src_pose"][b]) for b in range(B)])
                         .astype(np.float64)).to(dev)
    row("calc_flow_kernel (numpy rule, float64)", timed(lambda: prep.calc_flow_batch(
        x["depth_rendered"], x["depth_gt_observed"], T, synth.K_LINEMOD), reps, flush),
        20 * px * B, "8 B/px read + 12 B/px written; float64 arithmetic")
    pose_e = x["src_pose"]
    pose_g = torch.from_numpy(host["src_pose"] + np.float32(0.001)).to(dev)
    row("pose_error_kernel (ADD/ADI/ARP, N=3000)", timed(lambda: evaluation.pose_errors(
        pose_e, pose_g, x["point_cloud_model"], synth.K_LINEMOD), reps, flush), 3 * N * 4 * B,
        "ADI is an exact O(N^2) search: compute-bound (9e6 distances per pair), not HBM")
    vis_out = torch.empty_like(x["depth_rendered"])
    row("visib_mask_kernel", timed(lambda: evaluation.visib_mask(
        x["depth_gt_observed"], x["depth_rendered"], 0.015, out=vis_out), reps, flush), 12 * px * B,
        "two planes read, one written")
    row("flow_epe_kernel", timed(lambda: evaluation.flow_epe(
        step.zoom_flow, step.flow, step.valid, depth_rendered=x["depth_rendered"]), reps, flush),
        24 * px * B, "four flow planes + visible + depth read")

    # ---- reference gpu_flow_kernel.cu through its host-pointer _flow vs ours (8(d) item 5)
    ref = {}
    try:
        from oracle import flow_np

        if flow_np.reference_kernel_available():
            hs = [host["depth_rendered"], host["depth_gt_observed"], host["KT"], host["Kinv"]]
            for name, fn in (("reference _flow (host pointers, malloc + PCIe per call)",
                              lambda: flow_np.reference_kernel_gpu(*hs)),
                             ("ours _flow drop-in (rc_flow_host, staging arena)",
                              lambda: flow_mod.gpu_flow(*hs))):
                fn()
                t0 = time.perf_counter()
                for _ in range(5):
                    fn()
                ref[name] = {"ms_per_call": (time.perf_counter() - t0) / 5 * 1e3, "batch": B}
    except Exception as exc:  # the reference kernel is optional
        ref["error"] = repr(exc)

    print(json.dumps({"batch": B, "hbm_peak_GBps": pk, "kernels": rows, "host_pointer_flow": ref},
                     indent=1))


if __name__ == "__main__":
    main()
