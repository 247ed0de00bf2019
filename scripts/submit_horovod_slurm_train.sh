#!/bin/bash
#SBATCH --job-name=drn-horovod
#SBATCH --time=02:00:00
# Allreduce training under srun, one task per GPU/node (reference mkl-scripts/submit_horovod_cifar_cori_train.sh:
# per-rank batch = 128 / #ranks; rank and world size come from SLURM_PROCID / SLURM_NTASKS).
set -euo pipefail
HERE=$(cd "$(dirname "$0")/.." && pwd)
N=${SLURM_NTASKS:-${SLURM_NNODES:-1}}
BATCH=$(( 128 / N )); [ "$BATCH" -lt 1 ] && BATCH=1
export MASTER_ADDR=$(scontrol show hostnames "$SLURM_JOB_NODELIST" | head -1) MASTER_PORT=${MASTER_PORT:-29533}
srun -n "$N" ${PYTHON:-python3} "$HERE/resnet_cifar_main_horovod.py" --use_horovod=True --batch_size=$BATCH \
  --train_data_path="${CIFAR_DIR:-}" --log_root="$HERE/logs/horovod-$N/ckpt" --eval_dir="$HERE/logs/horovod-$N/train" \
  --num_gpus="${NUM_GPUS:-1}" --num_intra_threads="${INTRA:-0}" --num_inter_threads="${INTER:-0}" \
  --data_format="${DATA_FORMAT:-channels_first}" --train_steps="${STEPS:-80000}" --backend="${BACKEND:-torch}" "$@"
