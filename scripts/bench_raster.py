#!/usr/bin/env python
"""Rasteriser phase timings (CUDA events between the launches of one rc_render_dev call) for a
few scene shapes, L2 flushed between repetitions.

    python scripts/bench_raster.py [--batch 64] [--level 4] [--reps 20]

Cases: `offscreen` (every tile empty: the pure background-store rate of the fused raster
launch), `far` / `mid` / `near` (object at 1.2 / 0.9 / 0.5 m: few px per triangle ... many),
each with want = depth+image+mask (the refine loop's outputs) and depth only.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np
import torch

from mx_deepim_b200 import _lib, synth
from mx_deepim_b200.render import DeviceRenderer


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--level", type=int, default=4)
    ap.add_argument("--reps", type=int, default=20)
    a = ap.parse_args()
    B = a.batch
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(3)
    meshes = [synth.make_mesh(rng, level=a.level) for _ in range(4)]
    ren = DeviceRenderer(meshes, synth.K_LINEMOD, synth.W, synth.H, synth.ZNEAR, synth.ZFAR,
                         max_batch=B, device=dev)
    cls = torch.from_numpy(rng.integers(0, len(meshes), size=B).astype(np.int32)).to(dev)
    flush = torch.zeros(64 * 1024 * 1024, dtype=torch.float32, device=dev)
    out = {}
    res = {"batch": B, "triangles_per_mesh": len(meshes[0]["triangles"]), "cases": []}
    names = ["clear+setup+clip", "scan", "fill", "cover+resolve"]
    for case, z, off in (("offscreen", 1.0, 50.0), ("far", 1.2, 0.0), ("mid", 0.9, 0.0),
                         ("near", 0.5, 0.0)):
        poses = np.zeros((B, 3, 4), np.float32)
        for b in range(B):
            poses[b, :, :3] = synth.random_rotation(rng)
            poses[b, :, 3] = (off + rng.uniform(-0.05, 0.05), rng.uniform(-0.05, 0.05), z)
        pd = torch.from_numpy(poses).to(dev)
        for want in (("depth", "image", "mask"), ("depth",)):
            times = [[] for _ in names]
            for rep in range(a.reps + 3):
                flush.add_(1.0)
                marks = [_lib.Event() for _ in range(_lib.RC_RENDER_MARKS)]
                ren.render_batch(cls, pd, pixel_means=(1.0, 2.0, 3.0), want=want, out=out,
                                 marks=marks)
                torch.cuda.synchronize()
                if rep >= 3:
                    for j in range(len(names)):
                        times[j].append(marks[j].elapsed_ms(marks[j + 1]) * 1e3)
            med = {n: round(statistics.median(t), 2) for n, t in zip(names, times)}
            # the same call without marks: the passes are chained behind the background pass
            tot = []
            for rep in range(a.reps):
                flush.add_(1.0)
                e0, e1 = _lib.Event(), _lib.Event()
                e0.record()
                ren.render_batch(cls, pd, pixel_means=(1.0, 2.0, 3.0), want=want, out=out)
                e1.record()
                torch.cuda.synchronize()
                tot.append(e0.elapsed_ms(e1) * 1e3)
            med["serialised_total"] = round(sum(med.values()), 2)
            med["chained_total"] = round(statistics.median(tot), 2)
            st = ren.stats(B)
            fg = float((out["depth"][:B] > 0).sum().item())
            bytes_out = (4 + (12 if "image" in want else 0) + (4 if "mask" in want else 0)) * synth.H * synth.W * B
            res["cases"].append({"case": case, "want": "+".join(want), "us": med,
                                 "records": st["records"], "tile_entries": st["tile_entries"],
                                 "covered_px": fg,
                                 "render_GBps_out": round(bytes_out / med["chained_total"] / 1e3, 1)})
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
