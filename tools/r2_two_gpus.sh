#!/bin/bash
# Round 2, two GPUs: the sharded and in-process multi-GPU tests on real peers, then the driver's own bench command at N=2.
mkdir -p gpurun_out
set -o pipefail
timeout 900 python -m pytest tests/test_sharded.py tests/test_multi_store.py tests/test_store_semantics.py -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/pytest_2gpu.log
for lanes in 4 2; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus 2 --steps 20 --warmup 5 --lanes $lanes ${ALSO:-} 2> gpurun_out/bench_n2_l$lanes.err | tail -1 | tee gpurun_out/bench_n2_l$lanes.json | cut -c1-900
tail -3 gpurun_out/bench_n2_l$lanes.err
done
