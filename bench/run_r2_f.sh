#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests/test_epi2_gpu.py tests/test_engine_gpu.py -x -q > gpurun_out/r2f_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r2f_pytest.log
for v in "DRN_FUSE_BNBWD=0" "DRN_FUSE_BNBWD=1 DRN_FUSE_BNBWD_MINHW=56" "DRN_FUSE_BNBWD=1 DRN_FUSE_BNBWD_MINHW=28" "DRN_FUSE_BNBWD=1 DRN_FUSE_BNBWD_MINHW=14" "DRN_FUSE_BNBWD=1 DRN_FUSE_BNBWD_MINHW=7" "DRN_FUSE_BNBWD=1 DRN_FUSE_BNBWD_MINHW=14 DRN_NRES2=3"; do
  echo "== $v"; env $v python bench.py --steps 30 --warmup 5 --sustain_s 0 2>>gpurun_out/r2f_err.log | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'], d['clocks']['sm_mhz'], d['last_metrics'])"
done
tail -3 gpurun_out/r2f_err.log
