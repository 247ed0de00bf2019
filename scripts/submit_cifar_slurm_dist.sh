#!/bin/bash
#SBATCH --job-name=drn-cifar
#SBATCH --time=02:00:00
#SBATCH --nodes=9
#SBATCH --gres=gpu:1
# CIFAR-10 distributed run + evaluator on the last node (reference scripts/submit_cifar_daint_dist.sh,
# mkl-scripts/submit_ps_cifar_cori_dist.sh).
# usage: sbatch submit_cifar_slurm_dist.sh <#ps> <#workers> [batch per worker] [wipe]
set -euo pipefail
HERE=$(cd "$(dirname "$0")/.." && pwd)
PS=${1:-1}; WK=${2:-8}; BATCH=${3:-128}
RUN=$HERE/logs/$PS-ps-$WK-wk-batch-$BATCH-cifar10-log
[ -n "${4:-}" ] && rm -rf "$RUN"          # 4th argument: start from scratch instead of resuming
mkdir -p "$RUN"; cd "$RUN"
export TF_SCRIPT=$HERE/resnet_cifar_main.py TF_EVAL_SCRIPT=$HERE/resnet_cifar_eval.py TF_NUM_PS=$PS TF_NUM_WORKERS=$WK
export TF_FLAGS="--train_data_path=${CIFAR_DIR:-} --log_root=$RUN/ckpt --train_dir=$RUN/train --eval_dir=$RUN/train \
 --dataset=cifar10 --num_gpus=1 --batch_size=$BATCH --sync_replicas=True --train_steps=80000 --backend=${BACKEND:-torch} ${EXTRA_FLAGS:-}"
export TF_EVAL_FLAGS="--eval_data_path=${CIFAR_DIR:+$CIFAR_DIR/cifar-10-batches-bin/test_batch*} --log_root=$RUN/ckpt \
 --eval_dir=$RUN/eval --dataset=cifar10 --mode=eval --num_gpus=1"
exec "$HERE/scripts/run_dist_slurm.sh"
