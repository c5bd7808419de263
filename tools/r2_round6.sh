#!/bin/bash
mkdir -p gpurun_out
set -o pipefail
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/pytest.log
timeout 300 python tools/c5_mixed.py --seconds 10 --writers 2 --readers 8 2>&1 | tail -3 | tee gpurun_out/c5_one_gpu.log
timeout 300 python tools/c1_breakdown.py 2>&1 | tail -8 | tee gpurun_out/c1_breakdown.log
timeout 300 python tools/probe_ingest.py 2>&1 | tail -6 | tee gpurun_out/probe_ingest.log
