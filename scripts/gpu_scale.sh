#!/bin/bash
# N-GPU check on one box, the way the driver launches it: the default bench (cfg5, strong
# scaling) and the reference arm under torchrun.  Usage: bash scripts/gpu_scale.sh <tag> <N> [cfg4]
TAG=${1:-m}
N=${2:-2}
OUT=gpurun_out
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 400 $TR bench.py --gpus $N --steps 20 --warmup 4 > $OUT/bench_n${N}_$TAG.json 2> $OUT/bench_n${N}_$TAG.err; echo "rc=$?" >> $OUT/bench_n${N}_$TAG.err
timeout 200 $TR bench.py --impl reference --gpus $N --steps 2 --warmup 1 > $OUT/bench_ref_n${N}_$TAG.json 2> $OUT/bench_ref_n${N}_$TAG.err; echo "rc=$?" >> $OUT/bench_ref_n${N}_$TAG.err
if [ -n "$3" ]; then
timeout 300 $TR bench.py --gpus $N --config 4 --steps 20 --warmup 4 > $OUT/bench_cfg4_n${N}_$TAG.json 2> $OUT/bench_cfg4_n${N}_$TAG.err; echo "rc=$?" >> $OUT/bench_cfg4_n${N}_$TAG.err
fi
for f in bench_n${N}_$TAG bench_ref_n${N}_$TAG bench_cfg4_n${N}_$TAG; do [ -f $OUT/$f.json ] || continue; echo "== $f"; tail -2 $OUT/$f.err; python - <<PY
import json
try:
    d=json.loads(open("$OUT/$f.json").read().strip().splitlines()[-1])
    print("value %.0f ms/step %.4f n_gpus %s e2e %s limiter %s"%(d["value"], d["ms_per_step"], d["n_gpus"], d.get("e2e",{}).get("value"), d.get("e2e",{}).get("limiter")))
except Exception as e: print("failed", e)
PY
done
