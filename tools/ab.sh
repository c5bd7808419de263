#!/bin/bash
# A/B of tuning knobs on ONE box (boxes differ by a few % in HBM speed).
for rep in 1 2; do
for cfg in "" "TFP_DISABLE_LBEST=1" "TFP_REFRESH_NS=0"; do
  echo "== $cfg"
  env $cfg python tools/probe.py quick 2>&1 | grep -E "^C2|^d1024|k100"
done; done
