#!/bin/bash
# Evaluation process polling a checkpoint directory (reference scripts/submit_horovod_cifar_eval.sh,
# run_eval_cifar10_local.sh, run_local.sh).  usage: run_eval.sh cifar|imagenet <log_root> [flags...]
HERE=$(cd "$(dirname "$0")/.." && pwd)
KIND=${1:-cifar}; LOG_ROOT=${2:?log_root}; shift 2
if [ "$KIND" = imagenet ]; then S=resnet_imagenet_eval.py; D=imagenet; else S=resnet_cifar_eval.py; D=cifar10; fi
exec ${PYTHON:-python} "$HERE/$S" --dataset=$D --mode=eval --log_root="$LOG_ROOT" --eval_dir="$LOG_ROOT/../eval" \
  --num_gpus="${NUM_GPUS:-1}" "$@"
