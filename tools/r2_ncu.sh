#!/bin/bash
# Round 2 profiling visit (one GPU): the driver's bench command in full, then the recipe's two ncu passes on its
# headline part, then the tensor-core kernels of one 256-query batch.
mkdir -p gpurun_out
CMD="python bench.py --gpus 1 --steps 20 --warmup 5 --also none"
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 2> gpurun_out/bench_n1.err > gpurun_out/bench_n1.json; tail -2 gpurun_out/bench_n1.err; cut -c1-200 gpurun_out/bench_n1.json
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
timeout 300 $CMD > gpurun_out/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:scan_tma -s 40 -c 3 -o gpurun_out/prof_scan $CMD > gpurun_out/ncu_full.log 2>&1
timeout 200 python tools/mma_one.py > gpurun_out/mma_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mma_ -s 8 -c 4 -o gpurun_out/prof_mma python tools/mma_one.py > gpurun_out/ncu_mma.log 2>&1
tail -2 gpurun_out/ncu_full.log gpurun_out/ncu_mma.log
ls -la gpurun_out/*.ncu-rep
