#!/bin/bash
# Round 2, second GPU visit: the whole GPU suite with the in-process multi-GPU store, then the small-shard probe.
mkdir -p gpurun_out
set -o pipefail
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -25 | tee gpurun_out/pytest.log
timeout 300 python tools/probe_small.py 2>&1 | tail -1 | tee gpurun_out/probe_small2.log
