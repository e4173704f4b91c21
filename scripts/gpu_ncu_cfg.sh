#!/bin/bash
# ncu full capture of selected kernels of a refine-loop config.  Usage: bash scripts/gpu_ncu_cfg.sh <tag> <config> <kernel-regex> [skip] [count]
TAG=${1:-n}; CFG=${2:-3}; REGEX=${3:-raster_}; SKIP=${4:-30}; CNT=${5:-5}
OUT=gpurun_out
mkdir -p $OUT
CMD="python bench.py --config $CFG --steps 4 --warmup 3"
$CMD > $OUT/plain_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 $OUT/plain_$TAG.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"$REGEX" -s $SKIP -c $CNT -o $OUT/prof_$TAG -f $CMD > $OUT/ncu_f_$TAG.log 2>&1
tail -3 $OUT/ncu_f_$TAG.log
