#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python bench/prof_epi2.py > gpurun_out/r2g_epi2.log 2>&1; cat gpurun_out/r2g_epi2.log | tail -12
