#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2e_pytest.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/r2e_pytest.log
python bench/bench_input.py > gpurun_out/r2e_input.log 2>&1; echo "input rc=$?"; cat gpurun_out/r2e_input.log | tail -12
