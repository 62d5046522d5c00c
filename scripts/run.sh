#!/bin/bash
# Launcher: one process per GPU on this node (the counterpart of the reference's
# `srun -n 4 --mpi=pmix ./runme.sh`, README.md:16-18 / scripts/runme.sh).
#   scripts/run.sh <ngpus> <module-or-script> [args...]
# e.g. scripts/run.sh 8 rocm_mpi_b200.models.diffusion2d_ap --nx 65 --ny 128 --dimx 4 --dimy 2
# Several nodes (run the same line on every node): NNODES=2 NODE_RANK=<0|1> MASTER_ADDR=<node 0> scripts/run.sh 8 ...
# -- the halo exchange then defaults to the NCCL transport (peer mappings are CUDA IPC, i.e. node-local).
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
source "$HERE/setenv.sh"
N=${1:?number of GPUs}; shift
TARGET=${1:?python module or script}; shift
PORT=${MASTER_PORT:-29517}
if [[ "$TARGET" == *.py ]]; then RUN=("$TARGET"); else RUN=(-m "$TARGET"); fi
if [[ "$N" == "1" ]]; then
  exec python "${RUN[@]}" "$@"
fi
NNODES=${NNODES:-1}
if [[ "$NNODES" == "1" ]]; then
  exec python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port "$PORT" "${RUN[@]}" "$@"
fi
exec python -m torch.distributed.run --nnodes="$NNODES" --node-rank "${NODE_RANK:?NODE_RANK}" --nproc-per-node "$N" \
  --master-addr "${MASTER_ADDR:?MASTER_ADDR}" --master-port "$PORT" "${RUN[@]}" "$@"
