#!/bin/bash
# Round 2, fourth GPU visit: tensor-core hardening again, C5 tool functional run on one GPU, full suite
mkdir -p gpurun_out
set -o pipefail
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/pytest.log
timeout 600 python tools/probe_clustered.py 2>&1 | tail -4 | tee gpurun_out/probe_clustered.log
timeout 300 python tools/c5_mixed.py --seconds 10 --writers 2 --readers 8 2>&1 | tail -3 | tee gpurun_out/c5_one_gpu.log
