#!/bin/bash
# round-2 call B (2 GPUs): collectives + transports (push / pull / one-shot / NCCL arm)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests/test_comm_gpu.py -x -q -s > gpurun_out/r2b_comm.log 2>&1; echo "comm rc=$?"; tail -40 gpurun_out/r2b_comm.log
