#!/bin/bash
# ncu captures; raw-page CSV summaries are extracted on the box, only two full reports travel back
cd "$(dirname "$0")/.."
mkdir -p gpurun_out /tmp/ncu
NCU="ncu --set full --clock-control none --import-source on"
cap() { # name regex skip cmd...
  local name=$1 rx=$2 skip=$3; shift 3
  "$@" > /tmp/ncu/$name.plain.log 2>&1 && $NCU -k regex:$rx -s $skip -c 1 -f -o /tmp/ncu/$name "$@" > /dev/null 2>&1
  echo "$name rc=$?"
  ncu -i /tmp/ncu/$name.ncu-rep --page raw --csv > gpurun_out/ncu_$name.raw.csv 2>/dev/null
  ncu -i /tmp/ncu/$name.ncu-rep --page details --csv 2>/dev/null | grep -E "Duration|DRAM Throughput|Memory Throughput|Registers Per|Achieved Occupancy|Issue Slots Busy|L2 Hit|Warp Cycles Per Issued|No Eligible|Executed Ipc" > gpurun_out/ncu_$name.details.txt
}
cap wgrad_3x3_64_56 conv_wgrad 3 python bench/prof_one.py wgrad 128 56 64 64 3 1 0 0 3
cap wgrad_1x1_256_64_56 conv_wgrad 3 python bench/prof_one.py wgrad 128 56 256 64 1 1 0 0 3
cap wgrad_1x1_1024_256_14 conv_wgrad 3 python bench/prof_one.py wgrad 128 14 1024 256 1 1 0 0 3
cap bn_bwd_wide bn_bwd_finalize_apply 8 python bench/prof_bn2.py
cap igemm_bnbwd_stats_64_256_56 conv_igemm 3 python bench/prof_one.py fprop 128 56 64 256 1 1 2 1 3
cp /tmp/ncu/wgrad_3x3_64_56.ncu-rep /tmp/ncu/igemm_bnbwd_stats_64_256_56.ncu-rep gpurun_out/
ncu -i /tmp/ncu/igemm_bnbwd_stats_64_256_56.ncu-rep --page source --csv 2>/dev/null > gpurun_out/ncu_igemm_bnbwd_source.csv
ls -la gpurun_out
