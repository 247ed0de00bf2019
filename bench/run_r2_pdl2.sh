#!/bin/bash
# programmatic dependent launch with the LATE trigger: ImageNet and CIFAR steps per PDL mask
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
J='import sys,json; d=json.loads([l for l in sys.stdin if l.startswith("{")][-1]); print(d["value"], d["unit"], d["ms_per_step"], d["clocks"]["sm_mhz"])'
for m in 0 1 2 3; do
  echo "== DRN_PDL=$m"
  for i in 1 2; do DRN_PDL=$m timeout 300 python bench.py --steps 30 --warmup 5 --sustain_s 0 2>>gpurun_out/r2pdl2_err.log | python -c "$J"; done
  for b in 16 128; do DRN_PDL=$m timeout 300 python bench.py --model cifar10 --batch $b --steps 200 --warmup 20 --sustain_s 0 2>>gpurun_out/r2pdl2_err.log | python -c "$J"; done
done
timeout 600 python -m pytest tests/test_engine_gpu.py tests/test_parity_gpu.py tests/test_cli_gpu.py -m gpu -x -q > gpurun_out/r2pdl2_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2pdl2_pytest.log
