#!/bin/bash
# Quick GPU check: parity tests + default bench.  Usage: bash scripts/gpu_quick.sh <tag> [pytest -k expr]
TAG=${1:-q}
OUT=gpurun_out
mkdir -p $OUT
if [ -n "$2" ]; then
python -m pytest tests -m gpu -q -x -k "$2" > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_$TAG.log
else
python -m pytest tests -m gpu -q -x > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_$TAG.log
fi
python bench.py --steps 200 --warmup 10 --no-cpu-baseline > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench rc=$?" >> $OUT/bench_$TAG.err
tail -4 $OUT/pytest_$TAG.log; tail -2 $OUT/bench_$TAG.err; cat $OUT/bench_$TAG.json
