#!/bin/bash
#SBATCH --job-name=drn-cifar-single
#SBATCH --time=04:00:00
#SBATCH --nodes=2
#SBATCH --gres=gpu:1
# Serial trainer on the first node, evaluator on the last (reference scripts/submit_cifar_daint_single.sh).
set -euo pipefail
HERE=$(cd "$(dirname "$0")/.." && pwd)
RUN=$HERE/logs/single-cifar10-log; mkdir -p "$RUN"; cd "$RUN"
mapfile -t NODES < <(scontrol show hostnames "$SLURM_JOB_NODELIST")
srun --no-kill -n1 -N1 -w "${NODES[0]}" ${PYTHON:-python3} "$HERE/resnet_cifar_main.py" --train_data_path="${CIFAR_DIR:-}" \
  --log_root="$RUN/ckpt" --train_dir="$RUN/train" --eval_dir="$RUN/train" --dataset=cifar10 --num_gpus=1 \
  --batch_size=128 --train_steps=80000 > train.log 2>&1 &
T=$!
srun --no-kill -n1 -N1 -w "${NODES[-1]}" ${PYTHON:-python3} "$HERE/resnet_cifar_eval.py" \
  --eval_data_path="${CIFAR_DIR:+$CIFAR_DIR/cifar-10-batches-bin/test_batch*}" --log_root="$RUN/ckpt" --eval_dir="$RUN/eval" \
  --dataset=cifar10 --mode=eval --num_gpus=1 > eval.log 2>&1 &
E=$!
wait $T; sleep 70; kill $E 2>/dev/null || true
