#!/bin/bash
# round-2 call A: fresh-box test suite, headline bench, ncu captures of the kernels round 1 left without one
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2a_pytest.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench rc=$?"; cat gpurun_out/r2a_bench.json
python bench/prof_bn2.py > gpurun_out/r2a_bn.log 2>&1; tail -8 gpurun_out/r2a_bn.log
NCU="ncu --set full --clock-control none --import-source on"
python bench/prof_one.py wgrad 128 56 64 64 3 1 0 0 3 > gpurun_out/r2a_w1.log 2>&1 && \
  $NCU -k regex:conv_wgrad -s 3 -c 1 -f -o gpurun_out/ncu_wgrad_3x3_64_56 python bench/prof_one.py wgrad 128 56 64 64 3 1 0 0 3 > gpurun_out/r2a_w1n.log 2>&1
echo "ncu1 rc=$?"; cat gpurun_out/r2a_w1.log
python bench/prof_one.py wgrad 128 56 256 64 1 1 0 0 3 > gpurun_out/r2a_w2.log 2>&1 && \
  $NCU -k regex:conv_wgrad -s 3 -c 1 -f -o gpurun_out/ncu_wgrad_1x1_256_64_56 python bench/prof_one.py wgrad 128 56 256 64 1 1 0 0 3 > gpurun_out/r2a_w2n.log 2>&1
echo "ncu2 rc=$?"; cat gpurun_out/r2a_w2.log
python bench/prof_one.py wgrad 128 14 1024 256 1 1 0 0 3 > gpurun_out/r2a_w3.log 2>&1 && \
  $NCU -k regex:conv_wgrad -s 3 -c 1 -f -o gpurun_out/ncu_wgrad_1x1_1024_256_14 python bench/prof_one.py wgrad 128 14 1024 256 1 1 0 0 3 > gpurun_out/r2a_w3n.log 2>&1
echo "ncu3 rc=$?"; cat gpurun_out/r2a_w3.log
python bench/prof_bn2.py > /dev/null 2>&1 && \
  $NCU -k regex:bn_bwd_finalize_apply -s 8 -c 1 -f -o gpurun_out/ncu_bn_bwd_wide python bench/prof_bn2.py > gpurun_out/r2a_bnn.log 2>&1
echo "ncu4 rc=$?"
ls -la gpurun_out | tail -20
