#!/bin/bash
# Round 2, first GPU visit: the whole GPU suite, the driver's own bench command, small-shard stage sweep.
mkdir -p gpurun_out
set -o pipefail
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3 | tee gpurun_out/smoke.log
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 2> gpurun_out/bench_n1.err | tee gpurun_out/bench_n1.json | cut -c1-600
tail -5 gpurun_out/bench_n1.err
for st in 0 8 6 5 4; do
  TFP_SCAN_STAGES=$st timeout 300 python tools/probe_small.py 2>&1 | tail -1 | tee -a gpurun_out/probe_small.log
done
