#!/bin/bash
# L2 prefetch distance sweep for the tensor-core kernel (2M x 1024, 256 and 32 queries), then parity.
mkdir -p gpurun_out; rm -f gpurun_out/pf.log
for pf in 0 4 8 16 32; do
  echo "== TFP_MMA_PREFETCH=$pf" | tee -a gpurun_out/pf.log
  TFP_MMA_PREFETCH=$pf timeout 300 python - <<'PY' 2>&1 | tee -a gpurun_out/pf.log
import sys, os
sys.path.insert(0, 'tools')
import probe_mma
probe_mma.run(2_000_000, 1024, 256, 100, iters=3, label="q256")
probe_mma.run(2_000_000, 1024, 32, 100, iters=3, label="q32")
PY
done
timeout 600 python -m pytest tests/test_batched_mma.py -m gpu -x -q 2>&1 | tail -3 | tee -a gpurun_out/pf.log
