#!/bin/bash
# Allreduce ("Horovod") training: one process per GPU. The reference runs
# `mpirun -np 4 python3 resnet_cifar_main_horovod.py --use_horovod=True --batch_size=32`
# (scripts/submit-horovod-train-mac.sh); here torch.distributed.run plays mpirun's part.
# usage: submit_horovod_train.sh [nproc] [per-rank batch] [-- extra flags]
HERE=$(cd "$(dirname "$0")/.." && pwd)
NP=${1:-4}; BATCH=${2:-32}; shift $(( $# < 2 ? $# : 2 )); [ "${1:-}" = "--" ] && shift
exec ${PYTHON:-python} -m torch.distributed.run --nnodes=1 --nproc-per-node="$NP" --master-addr 127.0.0.1 \
  --master-port "${MASTER_PORT:-29511}" "$HERE/${TF_SCRIPT:-resnet_cifar_main_horovod.py}" --use_horovod=True \
  --batch_size="$BATCH" --num_gpus="${NUM_GPUS:-$NP}" --train_data_path="${CIFAR_DIR:-}" \
  --log_root="${LOG_ROOT:-$HERE/logs/horovod/ckpt}" --eval_dir="${LOG_ROOT:-$HERE/logs/horovod}/train" \
  --train_steps="${STEPS:-80000}" "$@"
