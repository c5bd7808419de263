#!/bin/bash
# N-GPU visit for the peer exchange: sharded tests, then C2 with the peer exchange and with ncclAllGather.
N=${1:-2}
mkdir -p gpurun_out
set -o pipefail
timeout 600 python -m pytest tests/test_sharded.py -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/p2p_pytest.log
for ex in p2p native; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --steps 400 --warmup 20 --exchange $ex 2> gpurun_out/p2p_bench_${ex}_$N.err | tail -1 | tee gpurun_out/p2p_bench_${ex}_$N.json
done
