#!/bin/bash
# Serial training on this machine (reference scripts/submit_mac_single.sh): no --job_name.
HERE=$(cd "$(dirname "$0")/.." && pwd)
exec ${PYTHON:-python} "$HERE/resnet_cifar_main.py" --train_data_path="${CIFAR_DIR:-}" --log_root="${LOG_ROOT:-$HERE/logs/single/ckpt}" \
  --train_dir="${LOG_ROOT:-$HERE/logs/single}/train" --eval_dir="${LOG_ROOT:-$HERE/logs/single}/train" \
  --dataset=cifar10 --num_gpus="${NUM_GPUS:-1}" --batch_size="${BATCH:-128}" --train_steps="${STEPS:-80000}" "$@"
