#!/bin/bash
# where does the tensor-core kernel's time go? (results are wrong with TFP_MMA_DEBUG set: timing only)
for dbg in 0 1 2 3; do
  echo "== TFP_MMA_DEBUG=$dbg"
  TFP_MMA_DEBUG=$dbg timeout 300 python - <<'PY'
import sys, os
sys.path.insert(0, 'tools')
import probe_mma
probe_mma.run(2_000_000, 1024, 256, 100, iters=3, label="q256")
probe_mma.run(2_000_000, 1024, 32, 100, iters=3, label="q32")
PY
done
