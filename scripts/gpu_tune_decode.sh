#!/bin/bash
# GPU box: rebuild with -D variants of the anchor-major decode and time it.  VARIANTS="-DYS_DEC_MINB=4|-DYS_DEC_GROUPS=1"
IFS='|' read -ra VS <<< "${VARIANTS:-}"
echo "== default"; python scripts/decode_only.py 20 2>&1 | tail -2
for v in "${VS[@]}"; do
  echo "== $v"
  YS_NVCC_EXTRA="$v" python yolo_scratch_b200/build.py --force > gpurun_out/tune_build.log 2>&1 || { tail -5 gpurun_out/tune_build.log; continue; }
  grep -A3 "k_decode_infer_amILb1" yolo_scratch_b200/build.log | grep -o "Used [0-9]* registers\|[0-9]* bytes spill stores" | tr '\n' ' '; echo
  python scripts/decode_only.py 20 2>&1 | tail -2
done
python yolo_scratch_b200/build.py --force > /dev/null 2>&1
