#!/bin/bash
# 8-GPU visit for the peer exchange: C2 with 2 and 4 finds in flight, then the north-star target shape.
mkdir -p gpurun_out
set -o pipefail
run() {  # name, extra args...
  name=$1; shift
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 \
      bench.py --gpus 8 --steps 400 --warmup 20 "$@" 2> gpurun_out/$name.err | tail -1 | tee gpurun_out/$name.json | cut -c1-400
}
run p2p8_c2_l2 --exchange p2p --lanes 2
run p2p8_c2_l4 --exchange p2p --lanes 4
run p2p8_target --exchange p2p --workload target --steps 200
