#!/bin/bash
# One-line profiling submission (reference mkl-scripts/profile_dist_ps_cori.sh: "sbatch -N 5
# submit_ps_cifar_cori_dist.sh 2 4 64 1"): 2 ps shards, 4 workers, 64 images per worker, start from
# scratch, with the parameter/FLOP table and the per-kernel timings of --profile in the worker logs.
HERE=$(cd "$(dirname "$0")" && pwd)
EXTRA_FLAGS="--profile=True ${EXTRA_FLAGS:-}" exec sbatch -N 5 "$HERE/submit_cifar_slurm_dist.sh" 2 4 64 1
