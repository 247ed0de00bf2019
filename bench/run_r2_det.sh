#!/bin/bash
# deterministic mode: tests, then what it costs (bench.py with DRN_DETERMINISTIC=1)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_deterministic_gpu.py -m gpu -x -q > gpurun_out/r2det_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r2det_pytest.log
for d in 0 1; do
  echo "DRN_DETERMINISTIC=$d"
  DRN_DETERMINISTIC=$d timeout 300 python bench.py --steps 20 --warmup 3 2>>gpurun_out/r2det_err.log | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'], d['clocks']['sm_mhz'], d['clocks']['reasons'])"
done
