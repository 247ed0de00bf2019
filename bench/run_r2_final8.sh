#!/bin/bash
# final 8-GPU check of the round: transports + deterministic data-parallel runs, then the three bench rows
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${N:-8}
timeout 500 python -m pytest tests/test_comm_gpu.py -m gpu -x -q -s -k all_gpus > gpurun_out/final8_comm_tests.log 2>&1; echo "comm tests rc=$?"; grep -c "steps ok" gpurun_out/final8_comm_tests.log; grep "deterministic\|passed\|failed" gpurun_out/final8_comm_tests.log
for mode in allreduce ps; do
  timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29813 bench.py --gpus $N --mode $mode --steps 30 --warmup 5 --sustain_s 0 >> gpurun_out/final8_bench.jsonl 2>> gpurun_out/final8_err.log; echo "$mode rc=$?"
done
cut -c1-260 gpurun_out/final8_bench.jsonl
