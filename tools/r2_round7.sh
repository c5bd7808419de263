#!/bin/bash
# deferred tensor-core status: the batched tests first, the probe with and without, C3 with one and two contexts, the suite
mkdir -p gpurun_out
set -o pipefail
timeout 600 python -m pytest tests/test_batched_mma.py -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/pytest_mma.log
timeout 200 python tools/probe_defer.py 2>&1 | tail -4 | tee gpurun_out/probe_defer.log
TFP_MMA_DEFER=0 timeout 200 python tools/probe_defer.py 2>&1 | tail -4 | tee -a gpurun_out/probe_defer.log
timeout 500 python bench.py --workload c3 --steps 20 --warmup 5 --also none > gpurun_out/c3_two_ctx.json 2> gpurun_out/c3_two_ctx.err; tail -c 1500 gpurun_out/c3_two_ctx.json
timeout 500 python bench.py --workload c3 --steps 20 --warmup 5 --also none --batch-lanes 1 > gpurun_out/c3_one_ctx.json 2> gpurun_out/c3_one_ctx.err; tail -c 600 gpurun_out/c3_one_ctx.json
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/pytest.log
