#!/bin/bash
# Where does the single-TF32 tensor-core kernel's time go?  (results are wrong with TFP_MMA_DEBUG set: timing only)
mkdir -p gpurun_out; rm -f gpurun_out/pair_diag.log
for dbg in 0 1 2 3; do
  echo "== TFP_MMA_DEBUG=$dbg" | tee -a gpurun_out/pair_diag.log
  TFP_MMA_DEBUG=$dbg timeout 300 python - <<'PY' 2>&1 | tee -a gpurun_out/pair_diag.log
import sys, os
sys.path.insert(0, 'tools')
import probe_mma
probe_mma.run(2_000_000, 1024, 256, 100, iters=3, label="q256")
probe_mma.run(2_000_000, 1024, 32, 100, iters=3, label="q32")
PY
done
timeout 200 python tools/mma_one.py > gpurun_out/mma_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mma_scan -s 3 -c 1 -o gpurun_out/prof_mma_pair2 \
    python tools/mma_one.py > gpurun_out/ncu_mma_pair2.log 2>&1
tail -2 gpurun_out/ncu_mma_pair2.log
