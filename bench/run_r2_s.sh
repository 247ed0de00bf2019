#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2s_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2s_pytest.log
python bench/probe_small_conv.py 2>&1 | head -7
for b in 128 16; do echo -n "cifar B=$b: "; python bench.py --model cifar10 --batch $b --steps 100 --warmup 10 --sustain_s 0 2>>gpurun_out/r2s_err.log | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'])"; done
python bench.py --steps 30 --warmup 5 --sustain_s 0 2>>gpurun_out/r2s_err.log | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('imagenet', d['value'], d['ms_per_step'])"
