#!/bin/bash
# Local multi-process launcher: builds the ps/worker topology on localhost and starts one process
# per task, like the reference's scripts/run_dist_tf_local.sh + submit_mac_dist.sh (ps on ports
# 2230+j, workers on 2220+i; ps tasks run without GPUs; logs ps<j>.log / wk<i>.log).
#
# usage: run_dist_local.sh <#ps> <#workers> [batch_size] [train_steps] [-- extra flags...]
# env:   TF_SCRIPT (default resnet_cifar_main.py)  TF_NUM_GPUS (GPUs per host, default: all visible)
#        TF_LOG_DIR (default ./logs/<ps>-ps-<wk>-wk-batch-<b>-log)  PYTHON (default python)
set -euo pipefail
NUM_PS=${1:-1}; NUM_WK=${2:-2}; BATCH=${3:-10}; STEPS=${4:-100}
shift $(( $# < 4 ? $# : 4 )) || true
[ "${1:-}" = "--" ] && shift
HERE=$(cd "$(dirname "$0")/.." && pwd)
PYTHON=${PYTHON:-python}
SCRIPT=${TF_SCRIPT:-resnet_cifar_main.py}
NGPU=${TF_NUM_GPUS:-$(${PYTHON} -c 'import torch; print(torch.cuda.device_count())' 2>/dev/null || echo 0)}
LOGDIR=${TF_LOG_DIR:-$HERE/logs/${NUM_PS}-ps-${NUM_WK}-wk-batch-${BATCH}-log}
mkdir -p "$LOGDIR"
ps_hosts=""; for ((j=0; j<NUM_PS; j++)); do ps_hosts+="${ps_hosts:+,}localhost:$((2230+j))"; done
wk_hosts=""; for ((i=0; i<NUM_WK; i++)); do wk_hosts+="${wk_hosts:+,}localhost:$((2220+i))"; done
COMMON="--ps_hosts=$ps_hosts --worker_hosts=$wk_hosts --sync_replicas=True --batch_size=$BATCH \
 --train_steps=$STEPS --num_gpus=$NGPU --log_root=$LOGDIR/ckpt --train_dir=$LOGDIR/train --eval_dir=$LOGDIR/train $*"
# A worker that loses a peer inside an in-kernel collective leaves with exit code 86 (watchdog ->
# CollectiveTimeout, train/loop.py); the launcher then stops the remaining tasks (their own PIDs) and
# starts the whole job again - it resumes from the latest checkpoint in $LOGDIR/ckpt, which is what
# MonitoredTrainingSession's session re-creation does for the reference. TF_MAX_RESTARTS (default 3).
MAX_RESTARTS=${TF_MAX_RESTARTS:-3}
attempt=0
while :; do
  pids=()
  for ((j=0; j<NUM_PS; j++)); do
    CUDA_VISIBLE_DEVICES= $PYTHON "$HERE/$SCRIPT" --job_name=ps --task_index=$j $COMMON >> "$LOGDIR/ps$j.log" 2>&1 &
    pids+=($!)
  done
  for ((i=0; i<NUM_WK; i++)); do
    $PYTHON "$HERE/$SCRIPT" --job_name=worker --task_index=$i $COMMON >> "$LOGDIR/wk$i.log" 2>&1 &
    pids+=($!)
  done
  echo "started ${#pids[@]} tasks (pids ${pids[*]}); logs in $LOGDIR"
  printf '%s\n' "${pids[@]}" > "$LOGDIR/pids"
  rc=0; for p in "${pids[@]}"; do wait "$p" || { r=$?; [ $rc -eq 86 ] || rc=$r; [ $r -eq 86 ] && for q in "${pids[@]}"; do kill "$q" 2>/dev/null || true; done; }; done
  if [ $rc -eq 86 ] && [ $attempt -lt $MAX_RESTARTS ]; then
    attempt=$((attempt+1)); echo "collective timeout (exit 86): restart $attempt/$MAX_RESTARTS from the latest checkpoint"
    rm -f "$LOGDIR/ckpt/train.done"; sleep 2; continue
  fi
  exit $rc
done
