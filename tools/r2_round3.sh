#!/bin/bash
# Round 2, third GPU visit: tensor-core path hardening (both filters), clustered-data probe
mkdir -p gpurun_out
set -o pipefail
timeout 1500 python -m pytest tests/test_batched_mma.py tests/test_find_parity.py -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/pytest_mma.log
TFP_MMA_TERMS=3 timeout 900 python -m pytest tests/test_batched_mma.py -m gpu -x -q 2>&1 | tail -5 | tee -a gpurun_out/pytest_mma.log
timeout 600 python tools/probe_clustered.py 2>&1 | tail -4 | tee gpurun_out/probe_clustered.log
