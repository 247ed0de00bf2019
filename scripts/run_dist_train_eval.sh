#!/bin/bash
# Training tasks plus a dedicated evaluator, the single-box analogue of the reference's
# scripts/run_dist_train_eval_daint.sh (which reserves the last node for the evaluator): the last
# GPU is kept for the eval process, the other GPUs host one worker (+ its ps shard) each.
# usage: run_dist_train_eval.sh <#ps> <#workers> [batch] [steps] [-- flags]
set -euo pipefail
HERE=$(cd "$(dirname "$0")" && pwd)
NUM_PS=${1:-1}; NUM_WK=${2:-2}; BATCH=${3:-128}; STEPS=${4:-80000}
LOGDIR=${TF_LOG_DIR:-$HERE/../logs/${NUM_PS}-ps-${NUM_WK}-wk-batch-${BATCH}-log}
export TF_LOG_DIR=$LOGDIR
"$HERE/run_dist_local.sh" "$@" &
TRAIN=$!
NG=$(${PYTHON:-python} -c 'import torch; print(torch.cuda.device_count())' 2>/dev/null || echo 0)
if [ "$NG" -gt 0 ]; then export CUDA_VISIBLE_DEVICES=$((NG-1)); fi
KIND=cifar; [ "${TF_SCRIPT:-}" = resnet_imagenet_main.py ] && KIND=imagenet
"$HERE/run_eval.sh" $KIND "$LOGDIR/ckpt" > "$LOGDIR/eval.log" 2>&1 &
EVAL=$!
wait $TRAIN; rc=$?
sleep "${EVAL_GRACE_SECS:-5}"; kill $EVAL 2>/dev/null || true
exit $rc
