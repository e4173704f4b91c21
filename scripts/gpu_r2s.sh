#!/bin/bash
# Short iteration: selected tests + default bench + the 128-pair share, serialised kernel times.
# Usage: bash scripts/gpu_r2s.sh <tag> "<pytest -k expr>"
TAG=${1:-s}
OUT=gpurun_out
mkdir -p $OUT
timeout 600 python -m pytest tests -m gpu -q --timeout 300 -k "$2" > $OUT/pytest_$TAG.log 2>&1; echo "rc=$?" >> $OUT/pytest_$TAG.log
timeout 400 python bench.py --steps 20 --warmup 4 --no-cpu-baseline > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "rc=$?" >> $OUT/bench_$TAG.err
timeout 300 python bench.py --pairs-total 128 --steps 20 --warmup 4 --no-cpu-baseline > $OUT/bench_b128_$TAG.json 2> $OUT/bench_b128_$TAG.err
CMD="python bench.py --steps 4 --warmup 4 --no-cpu-baseline --no-graph --min-seconds 0.01"
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"raster|frame|pm_|pose" -c 400 --csv --log-file $OUT/launches_$TAG.csv $CMD > $OUT/ncu_l_$TAG.log 2>&1
tail -3 $OUT/pytest_$TAG.log
python - <<PY
import json, csv, collections
for f in ("$TAG","b128_$TAG"):
    try:
        d=json.load(open("$OUT/bench_%s.json"%f)); print(f, "value %.0f ms/step %.4f e2e %.0f zoom_us %.1f"%(d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["kernel_us"])); print("   ", {k: round(v,1) for k,v in d.get("kernel_us_serialised").items()})
    except Exception as e: print(f, "failed", e)
try:
    rows=[r for r in csv.reader(open("$OUT/launches_$TAG.csv")) if len(r)>5]
    hdr=rows[0]; ki=hdr.index("Kernel Name"); vi=hdr.index("Metric Value"); ui=hdr.index("Metric Unit")
    agg=collections.defaultdict(list)
    for r in rows[1:]:
        try: v=float(r[vi].replace(",",""))
        except: continue
        if r[ui]=="ns": v/=1e3
        elif r[ui]=="ms": v*=1e3
        agg[r[ki].split("(")[0]].append(v)
    for k,v in sorted(agg.items(), key=lambda kv:-sum(kv[1])):
        print("%-40s n=%3d mean %8.1f us"%(k[:40],len(v),sum(v)/len(v)))
except Exception as e: print("launch list failed", e)
PY
