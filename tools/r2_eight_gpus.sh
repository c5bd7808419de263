#!/bin/bash
# Round 2, eight GPUs: the driver's own bench command at N=8 (headline + also.target / c3 / c4 / inproc), a lanes=2
# variant of the headline, and C5 (30 s) on one logical store over all eight GPUs in one process.
mkdir -p gpurun_out
set -o pipefail
N=${1:-8}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 \
    bench.py --gpus $N --steps 20 --warmup 5 2> gpurun_out/bench_n$N.err | tail -1 | tee gpurun_out/bench_n$N.json | cut -c1-700
tail -4 gpurun_out/bench_n$N.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 \
    bench.py --gpus $N --steps 20 --warmup 5 --lanes 2 --also none --inproc 0 2> gpurun_out/bench_n${N}_l2.err | tail -1 | tee gpurun_out/bench_n${N}_l2.json | cut -c1-400
timeout 400 python tools/c5_mixed.py --seconds 30 --writers 4 --readers 16 2>&1 | tail -2 | tee gpurun_out/c5_n$N.log
