#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -s > gpurun_out/r2i_pytest.log 2>&1; echo "pytest rc=$?"; grep -E "worst cosine|^native|^oracle|passed|failed" gpurun_out/r2i_pytest.log | tail -8
for b in 128 16; do echo -n "cifar B=$b: "; python bench.py --model cifar10 --batch $b --steps 100 --warmup 10 --sustain_s 0 2>>gpurun_out/r2i_err.log | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'])"; done
python bench.py --steps 20 --warmup 5 > gpurun_out/r2i_bench.json 2>>gpurun_out/r2i_err.log; cat gpurun_out/r2i_bench.json | cut -c1-300
tail -3 gpurun_out/r2i_err.log
