#!/bin/bash
# GPU box, round 2: full parity suite, smoke, both bench arms (default = cfg5), the smaller
# configs, the per-GPU share of an 8-GPU run, ncu launch list + full capture of the default bench.
# Usage: bash scripts/gpu_round2.sh <tag>
TAG=${1:-r2}
OUT=gpurun_out
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > $OUT/pytest_$TAG.log 2>&1; echo "rc=$?" >> $OUT/pytest_$TAG.log
timeout 300 python __graft_entry__.py smoke > $OUT/smoke_$TAG.log 2>&1; echo "smoke rc=$?" >> $OUT/smoke_$TAG.log
timeout 600 python bench.py --steps 20 --warmup 4 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "rc=$?" >> $OUT/bench_$TAG.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_ref_$TAG.json 2> $OUT/bench_ref_$TAG.err
for c in 2 3 4; do
timeout 300 python bench.py --config $c --steps 20 --warmup 4 --no-cpu-baseline > $OUT/bench_cfg${c}_$TAG.json 2> $OUT/bench_cfg${c}_$TAG.err
done
timeout 300 python bench.py --pairs-total 128 --steps 20 --warmup 4 --no-cpu-baseline > $OUT/bench_b128_$TAG.json 2> $OUT/bench_b128_$TAG.err
timeout 300 python bench.py --config 4 --pairs-total 32 --steps 20 --warmup 4 --no-cpu-baseline > $OUT/bench_cfg4_b32_$TAG.json 2> $OUT/bench_cfg4_b32_$TAG.err
tail -3 $OUT/pytest_$TAG.log; tail -2 $OUT/smoke_$TAG.log
python - <<PY
import json
for f in ("$TAG","cfg2_$TAG","cfg3_$TAG","cfg4_$TAG","b128_$TAG","cfg4_b32_$TAG"):
    try:
        d=json.load(open("$OUT/bench_%s.json"%f)); print(f, "value %.0f ms/step %.4f e2e %.0f zoom_us %.1f frac_comp %.3f"%(d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["kernel_us"], d["roofline"]["frac_compulsory"])); print("   ", d.get("kernel_us_serialised"))
    except Exception as e: print(f, "failed", e)
PY
