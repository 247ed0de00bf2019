#!/bin/bash
# round-2 call C (8 GPUs): collective tests at full width + the BASELINE configs 2-5 through bench.py
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${N:-8}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29700 + RANDOM % 200)) bench.py --gpus $N "$@" 2>> gpurun_out/r2c_err.log | tail -1 | tee -a gpurun_out/r2c_bench.jsonl; }
python -m pytest tests/test_comm_gpu.py -x -q -s > gpurun_out/r2c_comm_${N}gpu.log 2>&1; echo "comm rc=$?"; tail -30 gpurun_out/r2c_comm_${N}gpu.log
: > gpurun_out/r2c_bench.jsonl
run --steps 20 --warmup 5 --sustain_s 3
run --steps 20 --warmup 5 --sustain_s 0 --algo none
run --steps 20 --warmup 5 --sustain_s 0 --mode ps --pull push
run --steps 20 --warmup 5 --sustain_s 0 --mode ps --pull pull
run --steps 20 --warmup 5 --sustain_s 0 --batch 32
run --steps 20 --warmup 5 --sustain_s 0 --batch 32 --algo twoshot
run --steps 20 --warmup 5 --sustain_s 0 --batch 32 --algo none
run --steps 50 --warmup 10 --sustain_s 0 --model cifar10
run --steps 50 --warmup 10 --sustain_s 0 --model cifar10 --algo twoshot
run --steps 50 --warmup 10 --sustain_s 0 --model cifar10 --algo none
tail -5 gpurun_out/r2c_err.log
