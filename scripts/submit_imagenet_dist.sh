#!/bin/bash
# ImageNet ResNet-50 on one 8xB200 box, 8 ps shards x 8 workers, 128 images per worker (the
# reference's scripts/submit_imagenet_daint_dist.sh on 8 nodes): ps shard i and worker i share GPU i.
# usage: submit_imagenet_dist.sh [#ps] [#workers] [batch] [-- extra flags]
HERE=$(cd "$(dirname "$0")" && pwd)
PS=${1:-8}; WK=${2:-8}; BATCH=${3:-128}; shift $(( $# < 3 ? $# : 3 )); [ "${1:-}" = "--" ] && shift
TF_SCRIPT=resnet_imagenet_main.py exec "$HERE/run_dist_local.sh" "$PS" "$WK" "$BATCH" "${STEPS:-112600}" -- \
  --dataset=imagenet --train_data_path="${IMAGENET_DIR:-}" --num_epochs=90 "$@"
