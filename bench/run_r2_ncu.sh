#!/bin/bash
# one ncu capture per call: bench/run_r2_ncu.sh <name> <kernel regex> <command...>; summary extracted on the box
cd "$(dirname "$0")/.."
mkdir -p gpurun_out /tmp/ncu
name=$1; regex=$2; shift 2
"$@" > gpurun_out/ncu_${name}.plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/ncu_${name}.plain.log; exit 1; }
cat gpurun_out/ncu_${name}.plain.log | tail -2
ncu --set full --clock-control none --import-source on -k regex:"$regex" -s 3 -c 1 -f -o /tmp/ncu/$name "$@" > /dev/null 2>&1; echo "ncu rc=$?"
ncu -i /tmp/ncu/$name.ncu-rep --page details --csv 2>/dev/null | grep -E "Duration|DRAM Throughput|Memory Throughput|L2 Cache Throughput|Registers Per|Achieved Occupancy|Issue Slots Busy|Executed Ipc Active|Compute \(SM\) Throughput|Grid Size|Block Size" | cut -d, -f5,13- > gpurun_out/ncu_${name}.details_r2.txt
ncu -i /tmp/ncu/$name.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin)); hdr=rows[0]; val=rows[2] if len(rows)>2 else []; unit=rows[1]
want=['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active','smsp__issue_active.avg.pct_of_peak_sustained_active','smsp__inst_executed.sum','lts__t_sector_hit_rate.pct','l1tex__m_xbar2l1tex_read_bytes.sum','launch__registers_per_thread','launch__grid_size','smsp__average_warp_latency_per_inst_issued.ratio']
for k in want:
    if k in hdr: i=hdr.index(k); print(f'  {k} = {val[i]} {unit[i]}')
" > gpurun_out/ncu_${name}.raw_r2.txt
cat gpurun_out/ncu_${name}.raw_r2.txt
