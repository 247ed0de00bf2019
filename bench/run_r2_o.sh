#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests/test_parity_gpu.py -x -q -s > gpurun_out/r2o_pytest.log 2>&1; echo "pytest rc=$?"; grep -E "worst cosine|passed|failed|Error" gpurun_out/r2o_pytest.log | cut -c1-1500 | tail -5
