#!/bin/bash
# GPU box: rebuild the library with a list of -D variants and print the per-kernel times of each.
# usage: VARIANTS="-DYS_SCAN_THREADS=128|-DYS_RISK_MAX_CENTRES=8" bash scripts/gpu_tune.sh
mkdir -p gpurun_out
IFS='|' read -ra VS <<< "${VARIANTS:-}"
echo "== default"; python scripts/stage_times.py ${ARGS:-} 2>&1 | tail -2
for v in "${VS[@]}"; do
  echo "== $v"
  YS_NVCC_EXTRA="$v" python yolo_scratch_b200/build.py --force > gpurun_out/tune_build.log 2>&1 || { tail -5 gpurun_out/tune_build.log; continue; }
  python scripts/stage_times.py ${ARGS:-} 2>&1 | tail -2
done
