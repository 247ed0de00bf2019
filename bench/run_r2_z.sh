#!/bin/bash
# BatchNorm+ReLU in front of the last 1x1 convolution of a bottleneck fused into its prologue (DRN_FUSE_BN2 = max channels)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
DRN_FUSE_BN2=512 timeout 900 python -m pytest tests/test_engine_gpu.py -m gpu -x -q > gpurun_out/r2z_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2z_pytest.log
for c in 0 64 128 256 512; do
  echo "DRN_FUSE_BN2=$c"
  DRN_FUSE_BN2=$c timeout 300 python bench.py --steps 30 --warmup 5 2>>gpurun_out/r2z_err.log | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'], d['clocks']['sm_mhz'], d['clocks']['reasons'])"
done
