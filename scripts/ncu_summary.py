#!/usr/bin/env python
"""Turn the ncu artefacts a GPU round leaves in gpurun_out/ into the tracked summaries under
profiles/.

    python scripts/ncu_summary.py <tag>

reads   gpurun_out/launches_<tag>.csv   (ncu --metrics gpu__time_duration.sum launch list)
        gpurun_out/prof_<tag>.ncu-rep   (ncu --set full capture, one launch per kernel)
        gpurun_out/bench_<tag>.json     (the un-profiled bench line of the same command)
writes  profiles/<tag>_launches.csv     (id, kernel, grid, block, ns)
        profiles/<tag>_ncu_full.csv     (one row per captured kernel, selected raw metrics)
        profiles/<tag>_summary.md       (shares of the step, DRAM traffic vs algorithmic bytes)
"""
from __future__ import annotations

import csv
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
PROF = os.path.join(ROOT, "profiles")

METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__bytes_read.sum.pct_of_peak_sustained_elapsed",
    "dram__bytes_write.sum.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static",
    "launch__waves_per_multiprocessor", "launch__grid_size", "launch__block_size",
    "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__inst_executed.sum",
    "sm__inst_executed_pipe_lsu.sum", "smsp__cycles_active.avg",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
]

UNIT_SCALE = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "ns": 1e-3, "us": 1.0,
              "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}


def short(name):
    name = re.sub(r"\(.*", "", name)
    name = name.replace("void ", "").replace("rc::", "")
    return name.strip()


def launches(tag):
    path = os.path.join(OUT, "launches_%s.csv" % tag)
    if not os.path.exists(path):
        return []
    rows = []
    with open(path) as fh:
        lines = [l for l in fh if l.startswith('"')]
    rd = csv.reader(lines)
    hdr = next(rd)
    ik, ig, ib, iv, iu = (hdr.index(x) for x in ("Kernel Name", "Grid Size", "Block Size",
                                                   "Metric Value", "Metric Unit"))
    for r in rd:
        if len(r) <= iv:
            continue
        ns = float(r[iv].replace(",", "")) * (1.0 if r[iu] in ("ns", "nsecond") else 1e3)
        rows.append((int(r[0]), short(r[ik]), r[ig], r[ib], ns))
    return rows


def full(tag):
    rep = os.path.join(OUT, "prof_%s.ncu-rep" % tag)
    if not os.path.exists(rep):
        return [], []
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True,
                         text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    hdr, units = rows[0], rows[1]
    # this ncu prefixes some columns with their section ("FBSP.TriageCompute.dram__throughput..."):
    # match on the metric name itself
    where = {}
    for m in METRICS:
        for i, h in enumerate(hdr):
            if h == m or h.endswith("." + m):
                where[m] = i
                break
    cols = [m for m in METRICS if m in where]
    out = []
    for r in rows[2:]:
        d = {"kernel": short(r[hdr.index("Kernel Name")])}
        for m in cols:
            i = where[m]
            try:
                v = float(r[i].replace(",", ""))
            except ValueError:
                v = float("nan")
            u = units[i]
            if m.startswith("dram__bytes") or m.startswith("launch__shared"):
                v *= UNIT_SCALE.get(u, 1.0)
            elif m == "gpu__time_duration.sum":
                v *= UNIT_SCALE.get(u, 1.0)  # -> us
            d[m] = v
        out.append(d)
    return cols, out


def main():
    tag = sys.argv[1]
    os.makedirs(PROF, exist_ok=True)
    L = launches(tag)
    if L:
        with open(os.path.join(PROF, "%s_launches.csv" % tag), "w") as fh:
            fh.write("id,kernel,grid,block,ns\n")
            for r in L:
                fh.write('%d,%s,"%s","%s",%.0f\n' % r)
    cols, F = full(tag)
    if F:
        with open(os.path.join(PROF, "%s_ncu_full.csv" % tag), "w") as fh:
            fh.write("kernel," + ",".join(cols) + "\n")
            for d in F:
                fh.write(d["kernel"] + "," + ",".join("%.6g" % d[m] for m in cols) + "\n")
    bench = None
    bpath = os.path.join(OUT, "bench_%s.json" % tag)
    if os.path.exists(bpath):
        for line in open(bpath):
            if line.startswith("{"):
                bench = json.loads(line)
    md = ["# ncu summary `%s`" % tag, ""]
    if bench:
        md += ["Un-profiled bench line of the same workload (CUDA events, no profiler):", "",
               "```json", json.dumps({k: bench[k] for k in ("metric", "value", "unit", "ms_per_step",
                                                              "e2e", "roofline", "kernel_us", "clocks")
                                      if k in bench}, indent=1), "```", ""]
    if L:
        own = [r for r in L if not r[1].startswith("at::")]
        agg = {}
        for _, k, g, b, ns in own:
            a = agg.setdefault(k, [0, 0.0, g, b])
            a[0] += 1
            a[1] += ns
        tot = sum(a[1] for a in agg.values())
        md += ["## Launch list (`ncu --metrics gpu__time_duration.sum --clock-control none`)", "",
               "Serialised, cold-cache per-launch times; the SHARE column is what must agree with the",
               "CUDA-event breakdown of the bench line above.", "",
               "| kernel | launches | grid | block | mean us | share |", "|---|---|---|---|---|---|"]
        for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            md.append("| `%s` | %d | %s | %s | %.1f | %.1f %% |" % (k, a[0], a[2], a[3],
                                                                    a[1] / a[0] / 1e3, 100 * a[1] / tot))
        md.append("")
        if bench and "kernel_us" in bench:
            ku = bench["kernel_us"]
            t = sum(ku.values())
            md += ["CUDA-event shares of the same kernels inside the timed region: " +
                   ", ".join("%s %.1f %%" % (k, 100 * v / t) for k, v in ku.items()), ""]
    if F:
        md += ["## Full capture (`ncu --set full --clock-control none --import-source on`)", "",
               "| kernel | us | DRAM read MB | DRAM write MB | DRAM % peak | SM % peak | warps active % | regs | L2 hit % | smem bank conflicts | warp insts |",
               "|---|---|---|---|---|---|---|---|---|---|---|"]
        g = lambda d, m: d.get(m, float("nan"))
        for d in F:
            md.append("| `%s` | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f | %d | %.1f | %.3g | %.3g |" % (
                d["kernel"], g(d, "gpu__time_duration.sum"), g(d, "dram__bytes_read.sum") / 1e6,
                g(d, "dram__bytes_write.sum") / 1e6,
                # (this ncu leaves dram__throughput empty: read % + write % of the DRAM peak)
                g(d, "dram__throughput.avg.pct_of_peak_sustained_elapsed")
                if g(d, "dram__throughput.avg.pct_of_peak_sustained_elapsed") == g(d, "dram__throughput.avg.pct_of_peak_sustained_elapsed")
                else g(d, "dram__bytes_read.sum.pct_of_peak_sustained_elapsed") + g(d, "dram__bytes_write.sum.pct_of_peak_sustained_elapsed"),
                g(d, "sm__throughput.avg.pct_of_peak_sustained_elapsed"),
                g(d, "sm__warps_active.avg.pct_of_peak_sustained_active"),
                g(d, "launch__registers_per_thread"), g(d, "lts__t_sector_hit_rate.pct"),
                g(d, "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"),
                g(d, "smsp__inst_executed.sum")))
        md.append("")
    if F:
        # per-launch DRAM traffic of each captured kernel: what bench.py reports as roofline.traffic
        tpath = os.path.join(PROF, "dram_traffic.json")
        traffic = json.load(open(tpath)) if os.path.exists(tpath) else {}
        for d in F:
            name = re.sub(r"<.*", "", d["kernel"])
            rec = {
                "bytes_per_launch": g(d, "dram__bytes_read.sum") + g(d, "dram__bytes_write.sum"),
                "read": g(d, "dram__bytes_read.sum"), "write": g(d, "dram__bytes_write.sum"),
                "grid": g(d, "launch__grid_size"),
                "source": "profiles/%s_ncu_full.csv (ncu --set full, one launch)" % tag}
            traffic[name] = rec  # latest capture of the kernel, whatever its launch shape
            traffic["%s@%d" % (name, int(rec["grid"]))] = rec  # ... and per launch shape
        with open(tpath, "w") as fh:
            json.dump(traffic, fh, indent=1, sort_keys=True)
    with open(os.path.join(PROF, "%s_summary.md" % tag), "w") as fh:
        fh.write("\n".join(md) + "\n")
    print("\n".join(md))


if __name__ == "__main__":
    main()
