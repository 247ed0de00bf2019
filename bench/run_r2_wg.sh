#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
echo "== normal"; timeout 300 python bench/sweep_wgrad3.py
echo "== no red.global"; DRN_WGRAD_DEBUG=1 timeout 300 python bench/sweep_wgrad3.py
echo "== no MMA"; DRN_WGRAD_DEBUG=2 timeout 300 python bench/sweep_wgrad3.py
echo "== neither"; DRN_WGRAD_DEBUG=3 timeout 300 python bench/sweep_wgrad3.py
