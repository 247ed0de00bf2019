#!/bin/bash
# resident weight tile (DRN_B_RES): numerics, kernel timings, step A/B
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_conv_sweep_gpu.py tests/test_kernels_gpu.py tests/test_fused_entry_points_gpu.py tests/test_epi2_gpu.py tests/test_engine_gpu.py -m gpu -x -q > gpurun_out/r2bres_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2bres_pytest.log
for r in 0 1; do
  echo "== DRN_B_RES=$r"
  export DRN_B_RES=$r
  timeout 100 python bench/prof_one.py fprop 128 56 64 64 3 1 0 1 5
  timeout 100 python bench/prof_one.py fprop 128 56 64 64 3 1 2 1 5
  timeout 100 python bench/prof_one.py fprop 128 56 256 64 1 1 0 1 5
  timeout 100 python bench/prof_one.py fprop 128 56 64 256 1 1 1 1 5
  timeout 100 python bench/prof_one.py fprop 128 56 64 256 1 1 2 1 5
  timeout 100 python bench/prof_one.py fprop 128 32 16 16 3 1 0 1 5
  timeout 100 python bench/prof_one.py fprop 128 16 32 32 3 1 0 1 5
  timeout 100 python bench/prof_one.py fprop 128 8 64 64 3 1 0 1 5
  for i in 1 2; do timeout 300 python bench.py --steps 30 --warmup 5 --sustain_s 0 2>>gpurun_out/r2bres_err.log | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print(d['value'], d['unit'], d['ms_per_step'], d['clocks']['sm_mhz'])"; done
  for b in 16 128; do timeout 300 python bench.py --model cifar10 --batch $b --steps 200 --warmup 20 --sustain_s 0 2>>gpurun_out/r2bres_err.log | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('cifar', $b, d['value'], d['unit'], d['ms_per_step'])"; done
done
