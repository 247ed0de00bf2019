#!/bin/bash
# final 8-GPU confirmation: collective tests + the three headline transports + CIFAR
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${N:-8}
run() { timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29900 + RANDOM % 90)) bench.py --gpus $N "$@" 2>> gpurun_out/r2t_err.log | tail -1 | tee -a gpurun_out/r2t_bench.jsonl | cut -c1-260; }
python -m pytest tests/test_comm_gpu.py -x -q -s > gpurun_out/r2t_comm_${N}gpu.log 2>&1; echo "comm rc=$?"; tail -3 gpurun_out/r2t_comm_${N}gpu.log
: > gpurun_out/r2t_bench.jsonl
run --steps 20 --warmup 5 --sustain_s 0 --mode ps --pull push
run --steps 20 --warmup 5 --sustain_s 0
run --steps 50 --warmup 10 --sustain_s 0 --model cifar10
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29899 bench/bench_allreduce.py 2>/dev/null > gpurun_out/r2t_collectives_${N}gpu.jsonl; python -c "
import json
for l in open('gpurun_out/r2t_collectives_${N}gpu.jsonl'):
    if l.startswith('{'):
        d=json.loads(l); print(d['bytes'], d['blocks'], 'ps_push mm/p2p', d.get('ps_push_multimem_us'), d['ps_push_p2p_us'], 'twoshot', d.get('twoshot_multimem_us'))"
