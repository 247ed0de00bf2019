#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${N:-8}
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29811 bench/bench_allreduce.py > gpurun_out/r2p_collectives_${N}gpu.jsonl 2> gpurun_out/r2p_err.log; echo "collectives rc=$?"; cat gpurun_out/r2p_collectives_${N}gpu.jsonl | cut -c1-700
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29812 bench.py --gpus $N --impl torch_tuned --steps 20 --warmup 5 --sustain_s 0 > gpurun_out/r2p_tuned_${N}gpu.json 2>> gpurun_out/r2p_err.log; echo "tuned rc=$?"; cat gpurun_out/r2p_tuned_${N}gpu.json | cut -c1-400
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29813 bench.py --gpus $N --steps 20 --warmup 5 --sustain_s 3 > gpurun_out/r2p_native_${N}gpu.json 2>> gpurun_out/r2p_err.log; echo "native rc=$?"; cat gpurun_out/r2p_native_${N}gpu.json | cut -c1-300
tail -5 gpurun_out/r2p_err.log
