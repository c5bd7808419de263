#!/bin/bash
# 1-GPU visit for the tensor-core kernel: parity first, then the probe, then C3 (pairs on / off, 3xTF32).
mkdir -p gpurun_out
set -o pipefail
timeout 600 python -m pytest tests/test_batched_mma.py tests/test_find_parity.py -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/pair_pytest.log
timeout 300 python tools/probe_mma.py 2>&1 | tee gpurun_out/pair_probe_on.log
timeout 400 python bench.py --workload c3 --steps 20 --warmup 3 2>gpurun_out/c3.err | tee gpurun_out/c3_new.json | cut -c1-200
TFP_MMA_PAIR=0 timeout 400 python bench.py --workload c3 --steps 20 --warmup 3 2>>gpurun_out/c3.err | tee gpurun_out/c3_new_nopair.json | cut -c1-200
TFP_MMA_TERMS=3 timeout 400 python bench.py --workload c3 --steps 20 --warmup 3 2>>gpurun_out/c3.err | tee gpurun_out/c3_new_x3.json | cut -c1-200
