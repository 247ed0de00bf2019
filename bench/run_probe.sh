#!/bin/bash
# Runs bench/probe_conv.py case by case, restarting after a CUDA fault so one bad kernel
# does not hide the results of the others. Output: gpurun_out/probe.log
mkdir -p gpurun_out
LOG=gpurun_out/probe.log
: > $LOG
START=0
for attempt in $(seq 1 60); do
  timeout 300 python bench/probe_conv.py --start $START "$@" >> $LOG 2>&1
  rc=$?
  if [ $rc -eq 0 ] || [ $rc -eq 1 ]; then break; fi
  last=$(cat gpurun_out/probe_progress.txt 2>/dev/null || echo $START)
  echo "[run_probe] process exited rc=$rc at case $last; resuming at $((last+1))" >> $LOG
  START=$((last+1))
done
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv >> $LOG 2>&1
tail -100 $LOG
