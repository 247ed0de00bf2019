#!/bin/bash
# Slurm topology builder + per-node launcher: the counterpart of the reference's
# scripts/run_dist_tf_daint.sh / run_dist_train_eval_daint.sh.
#
# Inside an allocation it derives --ps_hosts / --worker_hosts from the node list (ps tasks on
# ports 2230+j, workers on 2220+i, ps co-located on worker nodes by default), writes one launcher
# script per node and starts it with `srun --no-kill -n1 -N1`, so that a failing node does not take
# the allocation down. ps tasks run without GPUs (CUDA_VISIBLE_DEVICES=); workers start after a
# short delay. With TF_EVAL_SCRIPT set, the LAST node of the allocation is reserved for the
# evaluator.
#
# Environment (same names as the reference): TF_SCRIPT TF_FLAGS TF_EVAL_SCRIPT TF_EVAL_FLAGS
# TF_NUM_PS TF_NUM_WORKERS TF_WORKER_PER_NODE TF_PS_PER_NODE TF_PS_IN_WORKER PYTHON
#
# Note: across nodes the trainers use the PyTorch backend's NCCL/gloo transports (--backend=torch);
# the in-kernel NVLink collectives of the native backend cover the GPUs of one NVSwitch box.
set -euo pipefail
: "${TF_SCRIPT:?TF_SCRIPT must name the training entry point}"
PYTHON=${PYTHON:-python3}
TF_FLAGS=${TF_FLAGS:-}
TF_NUM_PS=${TF_NUM_PS:-1}
TF_WORKER_PER_NODE=${TF_WORKER_PER_NODE:-1}
TF_PS_PER_NODE=${TF_PS_PER_NODE:-1}
TF_PS_IN_WORKER=${TF_PS_IN_WORKER:-true}
mapfile -t NODES < <(scontrol show hostnames "${SLURM_JOB_NODELIST:?not inside a Slurm allocation}")
NNODES=${#NODES[@]}
EVAL_NODE=""
if [ -n "${TF_EVAL_SCRIPT:-}" ]; then
  EVAL_NODE=${NODES[$((NNODES-1))]}
  NODES=("${NODES[@]:0:$((NNODES-1))}")
  NNODES=${#NODES[@]}
fi
TF_NUM_WORKERS=${TF_NUM_WORKERS:-$((NNODES * TF_WORKER_PER_NODE))}
JOB=${SLURM_JOB_ID:-local}

ps_hosts=""; declare -A PS_OF_NODE
for ((j=0; j<TF_NUM_PS; j++)); do
  if [ "$TF_PS_IN_WORKER" = true ]; then n=${NODES[$(( (j / TF_PS_PER_NODE) % NNODES ))]}
  else n=${NODES[$(( NNODES - 1 - (j / TF_PS_PER_NODE) ))]}; fi
  port=$((2230 + j % TF_PS_PER_NODE))
  ps_hosts+="${ps_hosts:+,}$n:$port"; PS_OF_NODE[$n]+=" $j"
done
wk_hosts=""; declare -A WK_OF_NODE
for ((i=0; i<TF_NUM_WORKERS; i++)); do
  n=${NODES[$(( (i / TF_WORKER_PER_NODE) % NNODES ))]}
  port=$((2220 + i % TF_WORKER_PER_NODE))
  wk_hosts+="${wk_hosts:+,}$n:$port"; WK_OF_NODE[$n]+=" $i"
done
echo "ps_hosts=$ps_hosts"; echo "worker_hosts=$wk_hosts"

pids=()
for n in "${NODES[@]}"; do
  s=".tfdist.$JOB.$n.sh"
  {
    echo "#!/bin/bash"
    for j in ${PS_OF_NODE[$n]:-}; do
      echo "CUDA_VISIBLE_DEVICES= $PYTHON $TF_SCRIPT $TF_FLAGS --ps_hosts=$ps_hosts --worker_hosts=$wk_hosts --job_name=ps --task_index=$j > ps.$JOB.$n-$j.log 2>&1 &"
    done
    [ -n "${PS_OF_NODE[$n]:-}" ] && echo "sleep 10"
    for i in ${WK_OF_NODE[$n]:-}; do
      echo "$PYTHON $TF_SCRIPT $TF_FLAGS --ps_hosts=$ps_hosts --worker_hosts=$wk_hosts --job_name=worker --task_index=$i > worker.$JOB.$n-$i.log 2>&1 &"
    done
    echo "wait"
  } > "$s"
  chmod +x "$s"
  srun --no-kill -n1 -N1 -w "$n" "./$s" &
  pids+=($!)
done
if [ -n "$EVAL_NODE" ]; then
  srun --no-kill -n1 -N1 -w "$EVAL_NODE" $PYTHON $TF_EVAL_SCRIPT ${TF_EVAL_FLAGS:-} > "eval.$JOB.$EVAL_NODE.log" 2>&1 &
  EVAL_PID=$!
fi
rc=0; for p in "${pids[@]}"; do wait "$p" || rc=$?; done
[ -n "${EVAL_PID:-}" ] && { sleep 70; kill "$EVAL_PID" 2>/dev/null || true; }
exit $rc
