#!/bin/bash
# Rasteriser iteration: raster/refine parity tests + cfg3/4/5 bench lines.  Usage: bash scripts/gpu_raster.sh <tag>
TAG=${1:-x}
OUT=gpurun_out
mkdir -p $OUT
python -m pytest tests -m gpu -q -x -k "raster or refine or step or render" > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_$TAG.log
tail -4 $OUT/pytest_$TAG.log
for c in 3 4 5; do
python bench.py --config $c --steps 40 --warmup 3 > $OUT/bench_cfg${c}_$TAG.json 2> $OUT/bench_cfg${c}_$TAG.err; echo "rc=$?" >> $OUT/bench_cfg${c}_$TAG.err
tail -1 $OUT/bench_cfg${c}_$TAG.err
python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_cfg${c}_$TAG.json"))
    print("cfg$c", round(d["value"]), "pair-it/s  ms/step", round(d["ms_per_step"],3), {k:round(v,1) for k,v in d["kernel_us_serialised"].items()})
except Exception as e: print("cfg$c failed", e)
PY
done
