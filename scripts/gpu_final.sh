#!/bin/bash
# GPU box, one GPU: the round's final evidence -- tests, smoke, the default bench line, the configs[4] line, the ncu launch
# list and one full capture of the loss-path kernels and the decode / NMS kernels.  Outputs in gpurun_out/.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_bench_reference_arm.json 2>> gpurun_out/final_bench.err; echo "ref arm rc=$?"
python bench.py --config c5 --steps 10 --no-nms > gpurun_out/final_bench_c5.json 2>> gpurun_out/final_bench.err; echo "c5 rc=$?"
CMD="python bench.py --steps 3 --warmup 3 --windows 2 --no-graph --no-cpu"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/final_launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"k_loss_stream|k_assign|k_pack_gt" -s 9 -c 3 -f -o gpurun_out/final_prof $CMD --no-nms > gpurun_out/ncu_full.log 2>&1
echo "full capture (loss path) rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"k_decode_infer|k_nms" -s 8 -c 6 -f -o gpurun_out/final_prof_aux $CMD > gpurun_out/ncu_full_aux.log 2>&1
echo "full capture (decode, nms) rc=$?"
python scripts/aux_kernel_times.py > gpurun_out/final_aux_kernel_times.json 2> gpurun_out/aux.err; echo "aux rc=$?"
ls -la gpurun_out | head -40
