#!/bin/bash
# ncu of one steady-state full pass of the tensor-core kernel on the C3 store (10M x 1024, 256 queries)
mkdir -p gpurun_out
timeout 400 python bench.py --workload c3 --steps 3 --warmup 3 > gpurun_out/c3_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mma_scan -s 7 -c 1 -o gpurun_out/prof_c3 \
    python bench.py --workload c3 --steps 3 --warmup 3 > gpurun_out/ncu_c3.log 2>&1
tail -2 gpurun_out/ncu_c3.log
