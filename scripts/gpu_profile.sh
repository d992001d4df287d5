#!/bin/bash
# Run on the GPU box: plain bench (must exit 0), then the ncu launch list and one --set full capture of the
# dominant kernel(s).  Outputs in gpurun_out/.
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --windows 2 --no-graph --no-cpu ${BENCH_EXTRA:---no-nms}"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"${KREGEX:-k_loss|k_assign}" -s 6 -c ${KCOUNT:-2} -f -o gpurun_out/prof $CMD > gpurun_out/ncu_full.log 2>&1
echo "full capture rc=$?"
ls -la gpurun_out
