#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2q_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2q_pytest.log
python bench/probe_small_conv.py > gpurun_out/r2q_small.log 2>&1; cat gpurun_out/r2q_small.log | tail -16
python bench.py --steps 30 --warmup 5 --sustain_s 3 2>>gpurun_out/r2q_err.log | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'], 'sustained', d['sustained']['value'], d['clocks'])"
