#!/bin/bash
# quick check of an epilogue change: conv numerics, three kernel timings, the step
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_conv_sweep_gpu.py tests/test_epi2_gpu.py tests/test_kernels_gpu.py -m gpu -x -q > gpurun_out/r2x_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r2x_pytest.log
timeout 120 python bench/prof_one.py fprop 128 56 64 256 1 1 2 1 5
timeout 120 python bench/prof_one.py fprop 128 56 64 256 1 1 1 1 5
timeout 120 python bench/prof_one.py fprop 128 56 64 64 3 1 0 1 5
for i in 1 2; do timeout 300 python bench.py --steps 30 --warmup 5 2>>gpurun_out/r2x_err.log | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'], d['clocks']['sm_mhz'], d['clocks']['reasons'])"; done
