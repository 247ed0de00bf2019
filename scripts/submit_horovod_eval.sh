#!/bin/bash
# One-shot CIFAR evaluation of a Horovod/allreduce run (reference scripts/submit_horovod_cifar_eval.sh:
# resnet_cifar_eval.py ... --eval_once=True).
# usage: submit_horovod_eval.sh <run dir> [data dir]
set -euo pipefail
HERE=$(cd "$(dirname "$0")/.." && pwd)
RUN=${1:?run directory (the one holding ckpt/)}; DATA=${2:-}
exec python "$HERE/resnet_cifar_eval.py" --eval_data_path="${DATA:+$DATA/cifar-10-batches-bin/test_batch*}" \
  --log_root="$RUN/ckpt" --eval_dir="$RUN/eval" --dataset=cifar10 --mode=eval --num_gpus=1 --eval_once=True
