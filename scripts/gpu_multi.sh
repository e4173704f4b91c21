#!/bin/bash
# Multi-GPU check on one box: default bench, reference arm and the sharded refine configs under
# torchrun, the way the driver launches them.  Usage: bash scripts/gpu_multi.sh <tag> <N>
TAG=${1:-m}
N=${2:-2}
OUT=gpurun_out
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
$TR bench.py --gpus $N --steps 200 --warmup 10 > $OUT/bench_n${N}_$TAG.json 2> $OUT/bench_n${N}_$TAG.err; echo "rc=$?" >> $OUT/bench_n${N}_$TAG.err
$TR bench.py --impl reference --gpus $N --steps 5 --warmup 1 > $OUT/bench_ref_n${N}_$TAG.json 2> $OUT/bench_ref_n${N}_$TAG.err; echo "rc=$?" >> $OUT/bench_ref_n${N}_$TAG.err
for c in 4 5; do
$TR bench.py --gpus $N --config $c --steps 40 --warmup 3 > $OUT/bench_cfg${c}_n${N}_$TAG.json 2> $OUT/bench_cfg${c}_n${N}_$TAG.err; echo "rc=$?" >> $OUT/bench_cfg${c}_n${N}_$TAG.err
done
python bench.py --steps 200 --warmup 10 --no-cpu-baseline > $OUT/bench_n1_$TAG.json 2> $OUT/bench_n1_$TAG.err
python -m pytest tests -m gpu -q -x -k "step or refine" > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_$TAG.log
tail -3 $OUT/pytest_$TAG.log
for f in bench_n1_$TAG bench_n${N}_$TAG bench_ref_n${N}_$TAG bench_cfg4_n${N}_$TAG bench_cfg5_n${N}_$TAG; do echo "== $f"; tail -2 $OUT/$f.err; cut -c1-900 $OUT/$f.json; done
