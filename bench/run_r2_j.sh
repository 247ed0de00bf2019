#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests/test_cli_gpu.py -x -q -k two_worker > gpurun_out/r2j_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/r2j_pytest.log
