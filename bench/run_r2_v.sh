#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out /tmp/ncu
python bench/bench_input.py > gpurun_out/r2v_input.log 2>&1; tail -9 gpurun_out/r2v_input.log
python bench/prof_comm_single.py > gpurun_out/r2v_comm1.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"allreduce_f32|ps_update" -s 8 -c 2 -f -o /tmp/ncu/comm1 python bench/prof_comm_single.py > /dev/null 2>&1; echo "ncu rc=$?"; cat gpurun_out/r2v_comm1.log | tail -3
ncu -i /tmp/ncu/comm1.ncu-rep --page raw --csv > gpurun_out/ncu_comm_single_rank.raw.csv 2>/dev/null
ncu -i /tmp/ncu/comm1.ncu-rep --page details --csv 2>/dev/null | grep -E "Duration|DRAM Throughput|Memory Throughput|Registers Per|Achieved Occupancy|Issue Slots Busy|Executed Ipc" | cut -d, -f5,13- > gpurun_out/ncu_comm_single_rank.details.txt
cat gpurun_out/ncu_comm_single_rank.details.txt | head -20
