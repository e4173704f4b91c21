#!/bin/bash
# ncu full capture of selected kernels of the default bench.  Usage: bash scripts/gpu_ncu.sh <tag> <kernel-regex> [bench args]
TAG=${1:-n}
REGEX=${2:-zoom_sample}
shift; shift
OUT=gpurun_out
mkdir -p $OUT
CMD="python bench.py --steps 6 --warmup 3 --no-cpu-baseline $@"
$CMD > $OUT/plain_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 $OUT/plain_$TAG.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"$REGEX" -s 8 -c 3 -o $OUT/prof_$TAG -f $CMD > $OUT/ncu_f_$TAG.log 2>&1
tail -3 $OUT/ncu_f_$TAG.log
