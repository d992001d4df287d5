#!/bin/bash
# GPU box: rebuild the library with a list of -D variants and print the per-kernel times of each.
# usage: VARIANTS="-DYS_SCAN_THREADS=128|-DYS_SCAN_STAGES=4 -DYS_SCAN_MINBLOCKS=3" bash scripts/gpu_tune.sh
mkdir -p gpurun_out
IFS='|' read -ra VS <<< "${VARIANTS:-}"
echo "== default"; python scripts/stage_times.py ${ARGS:-} 2>&1 | tail -1
for v in "${VS[@]}"; do
  echo "== $v"
  YS_NVCC_EXTRA="$v" python yolo_scratch_b200/build.py --force > gpurun_out/tune_build.log 2>&1 || { tail -5 gpurun_out/tune_build.log; continue; }
  grep -A2 "k_assignILb1\|k_loss_streamILb1ELi2" yolo_scratch_b200/build.log | grep -o "Used [0-9]* registers\|[0-9]* bytes spill stores" | tr '\n' ' '; echo
  python scripts/stage_times.py ${ARGS:-} 2>&1 | tail -1
done
