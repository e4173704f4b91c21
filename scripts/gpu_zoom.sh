#!/bin/bash
# GPU box: resampler iteration loop -- parity tests (per-test timeout), cfg2/cfg5 bench, one ncu capture.
# Usage: bash scripts/gpu_zoom.sh <tag> [ncu=1]
TAG=${1:-z}; NCU=${2:-1}
OUT=gpurun_out
mkdir -p $OUT
timeout 600 python -m pytest tests -m gpu -q -x --timeout 120 -k "zoom or step or golden or refine_gpu or smoke" > $OUT/pytest_$TAG.log 2>&1; echo "rc=$?" >> $OUT/pytest_$TAG.log; tail -3 $OUT/pytest_$TAG.log
for c in 2 5; do timeout 300 python bench.py --config $c --steps 20 --warmup 4 --no-cpu-baseline > $OUT/bench_cfg${c}_$TAG.json 2> $OUT/bench_cfg${c}_$TAG.err; done
python - <<PY
import json
for f in ("cfg2_$TAG","cfg5_$TAG"):
    try:
        d=json.load(open("$OUT/bench_%s.json"%f)); print(f, d["value"], d["ms_per_step"], d["roofline"]["kernel_us"], d["roofline"]["frac_compulsory"])
    except Exception as e: print(f, "failed", e)
PY
if [ "$NCU" = "1" ]; then
CMD="python bench.py --config 2 --steps 6 --warmup 3 --min-seconds 0 --no-cpu-baseline"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"zoom_sample" -s 6 -c 1 -o $OUT/prof_$TAG -f $CMD > $OUT/ncu_f_$TAG.log 2>&1
fi
