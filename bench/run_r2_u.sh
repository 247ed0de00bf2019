#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for v in 0 140 132 124 112; do echo -n "DRN_IGEMM_SMS=$v: "; DRN_IGEMM_SMS=$v python bench.py --steps 30 --warmup 5 --sustain_s 0 2>>gpurun_out/r2u_err.log | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'], d['clocks']['sm_mhz'])"; done
