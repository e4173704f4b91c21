#!/bin/bash
# GPU box: ncu launch list + full captures of the default bench (cfg5, eager issue so that ncu sees
# plain launches).  Usage: bash scripts/gpu_prof.sh <tag> [config] [kernel-regex]
TAG=${1:-p}; CFG=${2:-5}; REGEX=${3:-"zoom_sample|raster_|frame_tiles|pm_fwd|pm_bwd|pose_update|zoom_factor_boxes"}
OUT=gpurun_out
mkdir -p $OUT
CMD="python bench.py --config $CFG --steps 4 --warmup 4 --min-seconds 0 --no-cpu-baseline --no-graph"
$CMD > $OUT/plain_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 $OUT/plain_$TAG.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file $OUT/launches_$TAG.csv $CMD > $OUT/ncu_l_$TAG.log 2>&1
# one full capture per kernel of interest: skip the warm-up loops (launch-skip counts matching kernels only)
ncu --set full --clock-control none --import-source on -k regex:"$REGEX" -s 60 -c 16 -o $OUT/prof_$TAG -f $CMD > $OUT/ncu_f_$TAG.log 2>&1
tail -2 $OUT/ncu_f_$TAG.log; ls -la $OUT/prof_$TAG.ncu-rep
