#!/bin/bash
# size-based programmatic dependent launch on ImageNet ResNet-50: small layers only
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for cfg in "0 0" "400 0" "800 0" "1600 0" "800 400000" "1600 1700000" "0 1700000"; do
  set -- $cfg
  echo "DRN_PDL_SMALL_TILES=$1 DRN_PDL_SMALL_VECS=$2"
  DRN_PDL_SMALL_TILES=$1 DRN_PDL_SMALL_VECS=$2 timeout 300 python bench.py --steps 30 --warmup 5 2>>gpurun_out/r2pdl_err.log | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'], d['clocks']['sm_mhz'], d['clocks']['reasons'])"
done
