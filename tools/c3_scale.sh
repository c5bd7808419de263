#!/bin/bash
# C3 (10M x 1024, 256 queries, top-100) row-sharded over N GPUs, peer exchange; N=2 also re-runs the sharded tests
N=${1:-2}
mkdir -p gpurun_out
set -o pipefail
if [ "$N" == "2" ]; then
  timeout 600 python -m pytest tests/test_sharded.py -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/c3s_pytest.log
fi
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 \
    bench.py --gpus $N --workload c3 --steps 20 --warmup 3 2> gpurun_out/c3s_$N.err | tail -1 | tee gpurun_out/c3s_$N.json | cut -c1-300
