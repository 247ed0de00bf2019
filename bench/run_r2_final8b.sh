#!/bin/bash
# last 8-GPU rows of the round: ImageNet allreduce / parameter server, CIFAR 8 x 16
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${N:-8}
rm -f gpurun_out/final8b_bench.jsonl
run() { timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N "${@:2}" 2>> gpurun_out/final8b_err.log | grep '^{' >> gpurun_out/final8b_bench.jsonl; echo "rc=$? ${@:2}"; }
run 29821 --mode allreduce --steps 30 --warmup 5 --sustain_s 3
run 29822 --mode ps --steps 30 --warmup 5 --sustain_s 0
run 29823 --model cifar10 --steps 300 --warmup 30 --sustain_s 0
run 29824 --model cifar10 --algo none --steps 300 --warmup 30 --sustain_s 0
python - <<'PY'
import json
for l in open('gpurun_out/final8b_bench.jsonl'):
    d = json.loads(l)
    print(d['config'].get('model'), d['config'].get('parallelism'), d['value'], d['unit'], d['ms_per_step'], 'e2e', d.get('e2e', {}).get('value'), 'sustained', (d.get('sustained') or {}).get('value'), d['clocks']['sm_mhz'], d['clocks']['reasons'])
PY
