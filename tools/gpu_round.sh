#!/bin/bash
# One GPU-box visit: tests, smoke, probe, bench, then the two ncu passes of the profiling recipe.
mkdir -p gpurun_out
set -o pipefail
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -25 | tee gpurun_out/pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3 | tee gpurun_out/smoke.log
timeout 300 python tools/probe.py 2>&1 | tee gpurun_out/probe.log
timeout 600 python bench.py --steps 400 --warmup 20 2> gpurun_out/bench.err | tee gpurun_out/bench.json
if [ "$1" == "ncu" ]; then
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 20 --warmup 3 > gpurun_out/ncu_list.log 2>&1
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:scan_tma -s 10 -c 3 -o gpurun_out/prof_scan \
    python bench.py --steps 20 --warmup 3 > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
fi
