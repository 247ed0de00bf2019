#!/bin/bash
# Stops the tasks started by run_dist_local.sh for one log directory (the reference's
# scripts/kill.sh does `kill -9` on every python3 of the user; this one only touches recorded pids).
LOGDIR=${1:?usage: kill.sh <log dir written by run_dist_local.sh>}
[ -f "$LOGDIR/pids" ] || { echo "no pids file in $LOGDIR"; exit 1; }
while read -r p; do kill "$p" 2>/dev/null || true; done < "$LOGDIR/pids"
