#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2m_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r2m_pytest.log
for i in 1 2; do python bench.py --steps 30 --warmup 5 --sustain_s 3 2>>gpurun_out/r2m_err.log | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'], 'sustained', d['sustained']['value'], d['clocks']['sm_mhz'], d['clocks']['reasons'])"; done
python bench/prof_one.py fprop 128 56 64 256 1 1 2 1 5
python bench/prof_one.py fprop 128 56 64 256 1 1 1 1 5
python bench/prof_one.py fprop 128 56 64 64 3 1 0 1 5
