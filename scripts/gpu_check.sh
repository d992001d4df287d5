#!/bin/bash
# Run on the GPU box (through gpurun): build check, GPU parity tests (one process per file so a CUDA fault in one
# file does not poison the rest), smoke, and a short bench.  Logs go to gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
python __graft_entry__.py > gpurun_out/build.log 2>&1 || { echo "BUILD FAILED"; tail -20 gpurun_out/build.log; }
rc_all=0
for f in tests/test_gpu_assign.py tests/test_gpu_loss.py tests/test_gpu_nms.py tests/test_gpu_decode.py; do
  n=$(basename $f .py)
  timeout 900 python -m pytest $f -m gpu -q --tb=short -x > gpurun_out/$n.log 2>&1
  rc=$?; [ $rc -ne 0 ] && rc_all=$rc
  echo "== $n rc=$rc"; tail -${TAILN:-25} gpurun_out/$n.log
done
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "== smoke rc=$?"; tail -3 gpurun_out/smoke.log
timeout 900 python bench.py --steps ${STEPS:-20} --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "== bench rc=$?"; tail -c 3000 gpurun_out/bench.log; tail -5 gpurun_out/bench.err
exit $rc_all
