#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests/test_comm_gpu.py tests/test_cli_gpu.py -x -q -k "collectives or two_worker" > gpurun_out/r2r_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2r_pytest.log
N=2 bash -c 'timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29821 bench/bench_allreduce.py 2>/dev/null | cut -c1-500'
