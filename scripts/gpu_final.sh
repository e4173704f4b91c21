#!/bin/bash
# Short closing round on the B200 box: tests, smoke, both bench arms, the refine-loop configs and
# the ncu launch list + full capture of the default bench and cfg 3 (a subset of gpu_round.sh
# sized for a few GPU-minutes).  Usage (from the repo root, under gpurun): bash scripts/gpu_final.sh <tag>
TAG=${1:-r1}
OUT=gpurun_out
mkdir -p $OUT
timeout 120 python -m pytest tests -m gpu -q > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_$TAG.log
timeout 60 python __graft_entry__.py smoke > $OUT/smoke_$TAG.log 2>&1; echo "smoke rc=$?" >> $OUT/smoke_$TAG.log
timeout 90 python bench.py --impl reference --steps 10 --warmup 2 > $OUT/bench_ref_$TAG.json 2> $OUT/bench_ref_$TAG.err
timeout 150 python bench.py --steps 200 --warmup 10 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench rc=$?" >> $OUT/bench_$TAG.err
for c in 3 4 5; do
timeout 150 python bench.py --config $c --steps 40 --warmup 3 > $OUT/bench_cfg${c}_$TAG.json 2> $OUT/bench_cfg${c}_$TAG.err; echo "rc=$?" >> $OUT/bench_cfg${c}_$TAG.err
done
CMD="python bench.py --steps 6 --warmup 3 --no-cpu-baseline"
timeout 60 $CMD > $OUT/plain_$TAG.log 2>&1 && \
timeout 120 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/launches_$TAG.csv $CMD > $OUT/ncu_l_$TAG.log 2>&1
timeout 150 ncu --set full --clock-control none --import-source on -k regex:"zoom_sample|zoom_bbox|flow_kernel|pm_fwd" -s 12 -c 4 -o $OUT/prof_$TAG -f $CMD > $OUT/ncu_f_$TAG.log 2>&1
CMD3="python bench.py --config 3 --steps 4 --warmup 3 --no-graph"
timeout 60 $CMD3 > $OUT/plain3_$TAG.log 2>&1 && \
timeout 120 ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file $OUT/launches_cfg3_$TAG.csv $CMD3 > $OUT/ncu_l3_$TAG.log 2>&1
tail -3 $OUT/pytest_$TAG.log; tail -1 $OUT/smoke_$TAG.log; cat $OUT/bench_$TAG.json; cat $OUT/bench_ref_$TAG.json; cat $OUT/bench_cfg3_$TAG.json $OUT/bench_cfg4_$TAG.json $OUT/bench_cfg5_$TAG.json
