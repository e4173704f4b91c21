#!/bin/bash
# Full GPU round on the B200 box: tests, smoke, bench (both arms), refine-loop configs, per-kernel
# table, ncu launch lists + full captures (default bench and the refine loop).
# Usage (from the repo root, under gpurun):  bash scripts/gpu_round.sh <tag>
TAG=${1:-r1}
OUT=gpurun_out
mkdir -p $OUT
python -m pytest tests -m gpu -q > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_$TAG.log
python __graft_entry__.py smoke > $OUT/smoke_$TAG.log 2>&1; echo "smoke rc=$?" >> $OUT/smoke_$TAG.log
python bench.py --impl reference --steps 10 --warmup 2 > $OUT/bench_ref_$TAG.json 2> $OUT/bench_ref_$TAG.err
python bench.py --steps 200 --warmup 10 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench rc=$?" >> $OUT/bench_$TAG.err
for c in 3 4 5; do
python bench.py --config $c --steps 40 --warmup 3 > $OUT/bench_cfg${c}_$TAG.json 2> $OUT/bench_cfg${c}_$TAG.err; echo "rc=$?" >> $OUT/bench_cfg${c}_$TAG.err
done
python scripts/bench_kernels.py --batch 16 > $OUT/kernels_b16_$TAG.json 2> $OUT/kernels_b16_$TAG.err
python scripts/bench_kernels.py --batch 64 > $OUT/kernels_b64_$TAG.json 2> $OUT/kernels_b64_$TAG.err
python scripts/bench_raster.py > $OUT/raster_$TAG.json 2> $OUT/raster_$TAG.err
CMD="python bench.py --steps 6 --warmup 3 --no-cpu-baseline"
$CMD > $OUT/plain_$TAG.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/launches_$TAG.csv $CMD > $OUT/ncu_l_$TAG.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"zoom_sample|zoom_bbox|flow_kernel|pm_fwd" -s 12 -c 4 -o $OUT/prof_$TAG -f $CMD > $OUT/ncu_f_$TAG.log 2>&1
# refine-loop captures issue the launches eagerly (--no-graph) so that ncu sees plain launches
CMD3="python bench.py --config 3 --steps 4 --warmup 3 --no-graph"
$CMD3 > $OUT/plain3_$TAG.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file $OUT/launches_cfg3_$TAG.csv $CMD3 > $OUT/ncu_l3_$TAG.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"raster_|frame_tiles|zoom_factor_boxes" -s 60 -c 10 -o $OUT/prof_cfg3_$TAG -f $CMD3 > $OUT/ncu_f3_$TAG.log 2>&1
CMD5="python bench.py --config 5 --steps 4 --warmup 3 --no-graph"
$CMD5 > $OUT/plain5_$TAG.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file $OUT/launches_cfg5_$TAG.csv $CMD5 > $OUT/ncu_l5_$TAG.log 2>&1
tail -3 $OUT/pytest_$TAG.log; tail -1 $OUT/smoke_$TAG.log; cat $OUT/bench_$TAG.json; cat $OUT/bench_ref_$TAG.json; cat $OUT/bench_cfg3_$TAG.json $OUT/bench_cfg4_$TAG.json $OUT/bench_cfg5_$TAG.json
