#!/bin/bash
# product path vs bench e2e; input benchmark with concurrent decoders; ncu captures (wgrad x3, wide bn_bwd, two-side epilogue)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python resnet_imagenet_main.py --num_gpus=1 --batch_size=128 --train_steps=600 --log_every=100 --synthetic_data=True > gpurun_out/r2k_cli.log 2>&1; echo "cli rc=$?"; grep "images/sec" gpurun_out/r2k_cli.log | tail -4
python bench.py --steps 20 --warmup 5 > gpurun_out/r2k_bench.json 2>gpurun_out/r2k_bench.err; python -c "import json;d=json.load(open('gpurun_out/r2k_bench.json'));print('bench', d['value'], 'e2e', d['e2e']['value'], 'sustained', d['sustained']['value'], d['clocks'])"
python bench/bench_input.py > gpurun_out/r2k_input.log 2>&1; tail -6 gpurun_out/r2k_input.log
NCU="ncu --set full --clock-control none --import-source on"
python bench/prof_one.py wgrad 128 56 64 64 3 1 0 0 3 > gpurun_out/r2k_w1.log 2>&1 && $NCU -k regex:conv_wgrad -s 3 -c 1 -f -o gpurun_out/ncu_wgrad_3x3_64_56 python bench/prof_one.py wgrad 128 56 64 64 3 1 0 0 3 > /dev/null 2>&1; echo "ncu1 rc=$?"
python bench/prof_one.py wgrad 128 56 256 64 1 1 0 0 3 > gpurun_out/r2k_w2.log 2>&1 && $NCU -k regex:conv_wgrad -s 3 -c 1 -f -o gpurun_out/ncu_wgrad_1x1_256_64_56 python bench/prof_one.py wgrad 128 56 256 64 1 1 0 0 3 > /dev/null 2>&1; echo "ncu2 rc=$?"
python bench/prof_one.py wgrad 128 14 1024 256 1 1 0 0 3 > gpurun_out/r2k_w3.log 2>&1 && $NCU -k regex:conv_wgrad -s 3 -c 1 -f -o gpurun_out/ncu_wgrad_1x1_1024_256_14 python bench/prof_one.py wgrad 128 14 1024 256 1 1 0 0 3 > /dev/null 2>&1; echo "ncu3 rc=$?"
python bench/prof_bn2.py > gpurun_out/r2k_bn.log 2>&1 && $NCU -k regex:bn_bwd_finalize_apply -s 8 -c 1 -f -o gpurun_out/ncu_bn_bwd_wide python bench/prof_bn2.py > /dev/null 2>&1; echo "ncu4 rc=$?"
python bench/prof_one.py fprop 128 56 64 256 1 1 2 1 3 > gpurun_out/r2k_f1.log 2>&1 && $NCU -k regex:conv_igemm -s 3 -c 1 -f -o gpurun_out/ncu_igemm_bnbwd_stats_64_256_56 python bench/prof_one.py fprop 128 56 64 256 1 1 2 1 3 > /dev/null 2>&1; echo "ncu5 rc=$?"
cat gpurun_out/r2k_w*.log gpurun_out/r2k_f1.log
