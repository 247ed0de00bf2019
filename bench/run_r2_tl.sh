#!/bin/bash
# per-kernel table of one ImageNet ResNet-50 step (CUPTI timeline of the replayed graph)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python bench/timeline.py imagenet 128 > gpurun_out/r2tl_timeline.log 2>&1; echo "timeline rc=$?"; ls gpurun_out/timeline_* 2>/dev/null
python bench/step_table.py gpurun_out/timeline_imagenet_128.json > gpurun_out/step_table_imagenet_128_r2.txt; head -45 gpurun_out/step_table_imagenet_128_r2.txt | cut -c1-120
rm -f gpurun_out/timeline_*.json
