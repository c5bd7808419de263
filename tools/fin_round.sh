#!/bin/bash
# finalize kernel rework: parity under both filters, then the launch list of one batch
mkdir -p gpurun_out
set -o pipefail
timeout 600 python -m pytest tests/test_batched_mma.py tests/test_grpc_server.py -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/fin_pytest.log
TFP_MMA_TERMS=3 timeout 600 python -m pytest tests/test_batched_mma.py -m gpu -x -q 2>&1 | tail -3 | tee -a gpurun_out/fin_pytest.log
timeout 200 python tools/mma_one.py 2>&1 | tail -1 | tee gpurun_out/mma_plain.log &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/mma_launches.csv python tools/mma_one.py > gpurun_out/ncu_mma_list.log 2>&1
grep -E "mma_|qsplit" gpurun_out/mma_launches.csv | tail -4 | cut -d, -f5,15- | cut -c1-150
