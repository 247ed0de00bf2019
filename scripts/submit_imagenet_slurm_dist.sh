#!/bin/bash
#SBATCH --job-name=drn-imagenet
#SBATCH --time=24:00:00
#SBATCH --nodes=9
#SBATCH --gres=gpu:1
# ImageNet distributed run + evaluator (reference scripts/submit_imagenet_daint_dist.sh).
# usage: sbatch submit_imagenet_slurm_dist.sh <#ps> <#workers> [batch per worker] [wipe]
set -euo pipefail
HERE=$(cd "$(dirname "$0")/.." && pwd)
PS=${1:-8}; WK=${2:-8}; BATCH=${3:-128}
RUN=$HERE/logs/$PS-ps-$WK-wk-batch-$BATCH-imagenet-log
[ -n "${4:-}" ] && rm -rf "$RUN"
mkdir -p "$RUN"; cd "$RUN"
export TF_SCRIPT=$HERE/resnet_imagenet_main.py TF_EVAL_SCRIPT=$HERE/resnet_imagenet_eval.py TF_NUM_PS=$PS TF_NUM_WORKERS=$WK
export TF_FLAGS="--train_data_path=${IMAGENET_DIR:-} --log_root=$RUN/ckpt --train_dir=$RUN/train --dataset=imagenet \
 --num_gpus=1 --batch_size=$BATCH --sync_replicas=True --train_steps=112600 --num_epochs=90 --backend=${BACKEND:-torch} ${EXTRA_FLAGS:-}"
export TF_EVAL_FLAGS="--eval_data_path=${IMAGENET_DIR:-} --log_root=$RUN/ckpt --eval_dir=$RUN/eval --dataset=imagenet \
 --mode=eval --num_gpus=1 --num_epochs=3000"
exec "$HERE/scripts/run_dist_slurm.sh"
