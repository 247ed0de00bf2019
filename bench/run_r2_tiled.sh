#!/bin/bash
# tiled instead of im2col TMA loads for the A operand of 1x1 / stride-1 convolutions (DRN_A_TILED)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_conv_sweep_gpu.py tests/test_kernels_gpu.py tests/test_fused_entry_points_gpu.py tests/test_epi2_gpu.py tests/test_engine_gpu.py -m gpu -x -q > gpurun_out/r2tiled_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2tiled_pytest.log
for r in 0 1; do
  echo "== DRN_A_TILED=$r"
  export DRN_A_TILED=$r
  timeout 100 python bench/prof_one.py fprop 128 56 256 64 1 1 0 1 5
  timeout 100 python bench/prof_one.py fprop 128 56 64 256 1 1 1 1 5
  timeout 100 python bench/prof_one.py fprop 128 28 512 128 1 1 0 1 5
  timeout 100 python bench/prof_one.py fprop 128 14 1024 256 1 1 0 1 5
  timeout 100 python bench/prof_one.py fprop 128 7 512 2048 1 1 1 1 5
  timeout 100 python bench/prof_one.py wgrad 128 56 256 64 1 1 0 0 5
  timeout 100 python bench/prof_one.py wgrad 128 14 1024 256 1 1 0 0 5
  timeout 100 python bench/prof_one.py wgrad 128 7 512 2048 1 1 0 0 5
  for i in 1 2; do timeout 300 python bench.py --steps 30 --warmup 5 --sustain_s 0 2>>gpurun_out/r2tiled_err.log | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print(d['value'], d['unit'], d['ms_per_step'], d['clocks']['sm_mhz'])"; done
done
