#!/bin/bash
# Scaling lines the way the driver launches them.  Usage: bash scripts/gpu_scale.sh <tag> "<N list>"
TAG=${1:-s}
OUT=gpurun_out
mkdir -p $OUT
for N in ${2:-"8"}; do
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N"
if [ "$N" = "1" ]; then TR="python"; fi
$TR bench.py --gpus $N --steps 200 --warmup 10 --no-cpu-baseline > $OUT/bench_n${N}_$TAG.json 2> $OUT/bench_n${N}_$TAG.err; echo "rc=$?" >> $OUT/bench_n${N}_$TAG.err
for c in 4 5; do
$TR bench.py --gpus $N --config $c --steps 40 --warmup 3 > $OUT/bench_cfg${c}_n${N}_$TAG.json 2> $OUT/bench_cfg${c}_n${N}_$TAG.err; echo "rc=$?" >> $OUT/bench_cfg${c}_n${N}_$TAG.err
done
for f in bench_n${N}_$TAG bench_cfg4_n${N}_$TAG bench_cfg5_n${N}_$TAG; do tail -1 $OUT/$f.err; python - <<PY
import json
try:
    d=[json.loads(l) for l in open("$OUT/$f.json") if l.startswith("{")][-1]
    print("$f", d["n_gpus"], round(d["value"]), d["unit"], "ms/step", round(d["ms_per_step"],3), "e2e", round(d.get("e2e",{}).get("value",0)))
except Exception as e: print("$f failed", e)
PY
done; done
