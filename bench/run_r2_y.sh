#!/bin/bash
# ceilings: what would the step gain if a class of kernels were free (DRN_DEBUG_SKIP, wrong results)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for skip in "" wgrad bnbwd bnfwd wgrad,bnbwd,bnfwd; do
  echo "skip=[$skip]"
  DRN_DEBUG_SKIP=$skip timeout 300 python bench.py --steps 30 --warmup 5 2>>gpurun_out/r2y_err.log | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'], d['clocks']['sm_mhz'], d['clocks']['reasons'])"
done
