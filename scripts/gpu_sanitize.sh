#!/bin/bash
# compute-sanitizer passes over a small slice of the GPU parity tests (memcheck: out-of-bounds /
# misaligned accesses and leaks of the C-ABI library; racecheck: shared-memory hazards in the
# kernels that stage tiles, reduce through shared memory or keep a shared z-buffer).
# Usage (from the repo root, under gpurun):  bash scripts/gpu_sanitize.sh <tag> [per-pass timeout s]
TAG=${1:-san}
TMO=${2:-150}
OUT=gpurun_out
mkdir -p $OUT
SEL="tests/test_step_gpu.py::test_compare_step_matches_oracles tests/test_flow_gpu.py::test_flow_edge_cases tests/test_raster_gpu.py::test_near_plane_clipping_and_offscreen tests/test_pmloss_gpu.py::test_pm_loss_fused"
for tool in memcheck racecheck; do
  timeout $TMO compute-sanitizer --tool $tool --target-processes all --print-limit 20 \
    --log-file $OUT/sanitizer_${tool}_$TAG.log \
    python -m pytest $SEL -m gpu -q -x -p no:cacheprovider > $OUT/sanitizer_${tool}_pytest_$TAG.log 2>&1
  echo "$tool rc=$?" >> $OUT/sanitizer_${tool}_pytest_$TAG.log
  tail -2 $OUT/sanitizer_${tool}_pytest_$TAG.log
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY" $OUT/sanitizer_${tool}_$TAG.log | tail -3
done
