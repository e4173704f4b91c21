#!/bin/bash
# GPU box: smoke + both bench arms (default = cfg5) + the smaller configs.  Usage: bash scripts/gpu_bench.sh <tag> [steps]
TAG=${1:-b}; STEPS=${2:-20}
OUT=gpurun_out
mkdir -p $OUT
python __graft_entry__.py smoke > $OUT/smoke_$TAG.log 2>&1; echo "smoke rc=$?" >> $OUT/smoke_$TAG.log
python bench.py --steps $STEPS --warmup 4 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "rc=$?" >> $OUT/bench_$TAG.err
python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_ref_$TAG.json 2> $OUT/bench_ref_$TAG.err
for c in 2 3 4; do
python bench.py --config $c --steps $STEPS --warmup 4 --no-cpu-baseline > $OUT/bench_cfg${c}_$TAG.json 2> $OUT/bench_cfg${c}_$TAG.err; echo "rc=$?" >> $OUT/bench_cfg${c}_$TAG.err
done
tail -3 $OUT/smoke_$TAG.log; tail -3 $OUT/bench_$TAG.err; cut -c1-1500 $OUT/bench_$TAG.json
