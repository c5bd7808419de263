#!/bin/bash
# The profiling recipe's two passes on bench.py, plus the tensor-core kernel (both filters).
mkdir -p gpurun_out
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 20 --warmup 3 > gpurun_out/ncu_list.log 2>&1
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:scan_tma -s 10 -c 3 -o gpurun_out/prof_scan \
    python bench.py --steps 20 --warmup 3 > gpurun_out/ncu_full.log 2>&1
timeout 200 python tools/mma_one.py > gpurun_out/mma_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mma_scan -s 3 -c 1 -o gpurun_out/prof_mma1 \
    python tools/mma_one.py > gpurun_out/ncu_mma1.log 2>&1
TFP_MMA_TERMS=3 timeout 200 python tools/mma_one.py > gpurun_out/mma_plain3.log 2>&1 &&
TFP_MMA_TERMS=3 timeout 600 ncu --set full --clock-control none --import-source on -k regex:mma_scan -s 3 -c 1 -o gpurun_out/prof_mma3 \
    python tools/mma_one.py > gpurun_out/ncu_mma3.log 2>&1
tail -2 gpurun_out/ncu_full.log gpurun_out/ncu_mma1.log gpurun_out/ncu_mma3.log
