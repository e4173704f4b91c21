#!/bin/bash
# Round-2 iteration check on the B200 box: full parity suite (all failures, not -x), the default
# bench (cfg5), the per-GPU share of an 8-GPU run, cfg3 / cfg4, and a pure-write / copy bandwidth
# probe for the resampler's roofline.  Usage: bash scripts/gpu_r2q.sh <tag> [skip-tests]
TAG=${1:-q}
OUT=gpurun_out
mkdir -p $OUT
if [ -z "$2" ]; then
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > $OUT/pytest_$TAG.log 2>&1; echo "rc=$?" >> $OUT/pytest_$TAG.log
fi
timeout 400 python bench.py --steps 20 --warmup 4 --no-cpu-baseline > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "rc=$?" >> $OUT/bench_$TAG.err
timeout 300 python bench.py --pairs-total 128 --steps 20 --warmup 4 --no-cpu-baseline > $OUT/bench_b128_$TAG.json 2> $OUT/bench_b128_$TAG.err
for c in 3 4; do
timeout 300 python bench.py --config $c --steps 20 --warmup 4 --no-cpu-baseline > $OUT/bench_cfg${c}_$TAG.json 2> $OUT/bench_cfg${c}_$TAG.err
done
timeout 120 python - > $OUT/bw_$TAG.json 2> $OUT/bw_$TAG.err <<PY
import json, torch
n = 4 << 30
a = torch.empty(n, dtype=torch.uint8, device="cuda"); b = torch.empty(n, dtype=torch.uint8, device="cuda")
def t(f, reps=10):
    f(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); f(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best
w = t(lambda: a.zero_()); c = t(lambda: b.copy_(a)); r = t(lambda: a.view(torch.int32).sum())
print(json.dumps({"write_gbs": n / w / 1e6, "copy_gbs": 2 * n / c / 1e6, "read_gbs": n / r / 1e6}))
PY
tail -5 $OUT/pytest_$TAG.log; cat $OUT/bw_$TAG.json
python - <<PY
import json
for f in ("$TAG","b128_$TAG","cfg3_$TAG","cfg4_$TAG"):
    try:
        d=json.load(open("$OUT/bench_%s.json"%f)); print(f, "value %.0f ms/step %.4f e2e %.0f zoom_us %.1f frac_comp %.3f"%(d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["kernel_us"], d["roofline"]["frac_compulsory"])); print("   ", d.get("kernel_us_serialised")); print("   ", d.get("rasteriser"))
    except Exception as e: print(f, "failed", e)
PY
