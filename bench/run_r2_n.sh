#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python bench/timeline.py cifar10 16 > /dev/null 2>&1; python bench/step_table.py gpurun_out/timeline_cifar10_16.json > gpurun_out/r2n_cifar16.txt; head -30 gpurun_out/r2n_cifar16.txt
python bench/timeline.py cifar10 128 > /dev/null 2>&1; python bench/step_table.py gpurun_out/timeline_cifar10_128.json > gpurun_out/r2n_cifar128.txt; head -30 gpurun_out/r2n_cifar128.txt
python bench/step_table.py gpurun_out/timeline_cifar10_16.json augment all > gpurun_out/r2n_cifar16_events.txt
rm -f gpurun_out/timeline_*.json
