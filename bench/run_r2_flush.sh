#!/bin/bash
# barrier-free statistics flush: numerics, small-kernel fixed cost, CIFAR and ImageNet steps
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_conv_sweep_gpu.py tests/test_kernels_gpu.py tests/test_fused_entry_points_gpu.py tests/test_epi2_gpu.py tests/test_engine_gpu.py tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/r2flush_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2flush_pytest.log
timeout 200 python bench/probe_small_conv.py 2>&1 | grep "N= 16"
for b in 16 128; do timeout 300 python bench.py --model cifar10 --batch $b --steps 200 --warmup 20 --sustain_s 0 2>>gpurun_out/r2flush_err.log | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('cifar', $b, d['value'], d['unit'], d['ms_per_step'])"; done
for i in 1 2; do timeout 300 python bench.py --steps 30 --warmup 5 --sustain_s 0 2>>gpurun_out/r2flush_err.log | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print(d['value'], d['unit'], d['ms_per_step'], d['clocks']['sm_mhz'])"; done
