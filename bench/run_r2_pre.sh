#!/bin/bash
# generalised BatchNorm prologue (any filter size, 16/32-channel tiles) + finalize by the last CTA:
# kernel tests, engine oracle tests with everything fused, CIFAR and ImageNet A/B
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_fused_entry_points_gpu.py -m gpu -x -q -k "prologue or finalize_by" > gpurun_out/r2pre_kernels.log 2>&1; echo "kernel tests rc=$?"; tail -4 gpurun_out/r2pre_kernels.log
DRN_FUSE_PRE=all timeout 900 python -m pytest tests/test_engine_gpu.py tests/test_deterministic_gpu.py -m gpu -x -q > gpurun_out/r2pre_engine_all.log 2>&1; echo "engine tests (all fused) rc=$?"; tail -4 gpurun_out/r2pre_engine_all.log
timeout 900 python -m pytest tests/test_engine_gpu.py tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/r2pre_engine_default.log 2>&1; echo "engine tests (default) rc=$?"; tail -4 gpurun_out/r2pre_engine_default.log
run() { python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print(d['value'], d['unit'], d['ms_per_step'], d['clocks']['sm_mhz'])"; }
for b in 16 128; do
  for cfg in "none 1" "default 0" "default 1"; do
    set -- $cfg
    echo "cifar10 batch $b DRN_FUSE_PRE=$1 DRN_FIN_IN_CONV=$2"
    DRN_FUSE_PRE=$1 DRN_FIN_IN_CONV=$2 timeout 300 python bench.py --model cifar10 --batch $b --steps 200 --warmup 20 --sustain_s 0 2>>gpurun_out/r2pre_err.log | run
  done
done
for cfg in "default 1" "all 1"; do
  set -- $cfg
  echo "imagenet DRN_FUSE_PRE=$1 DRN_FIN_IN_CONV=$2"
  DRN_FUSE_PRE=$1 DRN_FIN_IN_CONV=$2 timeout 300 python bench.py --steps 30 --warmup 5 --sustain_s 0 2>>gpurun_out/r2pre_err.log | run
done
