"""Host-pointer GT flow (the reference's `_flow` / gpu_flow contract: pageable numpy in and out)
timed per call at B = 16: the reference's own gpu_flow_kernel.cu (oracle/_ref, unmodified), our
plain pageable-copy path (RC_FLOW_HOST_THREADS=1) and the pipelined path with 2 / 4 / 8 workers.
Every variant's result is compared bit for bit with the first one.

    python scripts/bench_flow_host.py [--batch 16] [--reps 8]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--reps", type=int, default=8)
    args = ap.parse_args()
    from mx_deepim_b200 import flow as flow_mod
    from mx_deepim_b200 import synth
    from oracle import flow_np

    host = synth.make_compare_batch(args.batch, 20181804 + 2)
    hs = [host["depth_rendered"], host["depth_gt_observed"], host["KT"], host["Kinv"]]
    out, base = {}, None

    def timed(name, fn):
        nonlocal base
        f, v = fn()
        if base is None:
            base = (f, v)
        same = bool(np.array_equal(f, base[0]) and np.array_equal(v, base[1]))
        t0 = time.perf_counter()
        for _ in range(args.reps):
            fn()
        out[name] = {"ms_per_call": (time.perf_counter() - t0) / args.reps * 1e3, "bit_identical": same}

    for n in (1, 2, 4, 8):
        os.environ["RC_FLOW_HOST_THREADS"] = str(n)
        timed("ours rc_flow_host, %d worker%s" % (n, "" if n == 1 else "s (pipelined)"),
              lambda: flow_mod.gpu_flow(*hs))
    os.environ.pop("RC_FLOW_HOST_THREADS")
    timed("ours rc_flow_host, default", lambda: flow_mod.gpu_flow(*hs))
    if flow_np.reference_kernel_available():
        timed("reference _flow (its own .cu, malloc + pageable copies per call)",
              lambda: flow_np.reference_kernel_gpu(*hs))
    print(json.dumps({"batch": args.batch, "host_pointer_flow": out}, indent=1))


if __name__ == "__main__":
    main()
