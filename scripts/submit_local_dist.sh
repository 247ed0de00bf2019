#!/bin/bash
# 1 ps + 2 workers on localhost, CPU only, batch 10, 100 steps: the reference's local smoke run
# (scripts/submit_mac_dist.sh; README calls it submit_local_dist.sh).
HERE=$(cd "$(dirname "$0")" && pwd)
TF_NUM_GPUS=0 exec "$HERE/run_dist_local.sh" 1 2 "${1:-10}" "${2:-100}" -- --resnet_size=${RESNET_SIZE:-50} "${@:3}"
