#!/bin/bash
# Environment for a run (the counterpart of the reference's scripts/setenv.sh, which loads ROCm / MPI
# modules and picks ROCm-aware vs standard MPI with IGG_ROCMAWARE_MPI, :11-18).
REPO="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
export PYTHONPATH="$REPO:${PYTHONPATH:-}"
export MASTER_ADDR=${MASTER_ADDR:-127.0.0.1}
export OMP_NUM_THREADS=${OMP_NUM_THREADS:-1}

# Halo transport: peer (NVLink peer stores + flags; default), nccl (pack / NCCL p2p / unpack baseline),
# host (pinned staging + gloo: the IGG_ROCMAWARE_MPI=0 analogue).
export IGG_B200_TRANSPORT=${IGG_B200_TRANSPORT:-peer}
# Per dimension (upstream: IGG_ROCMAWARE_MPI_DIMX/_DIMY/_DIMZ): with transport nccl or host,
# IGG_B200_TRANSPORT_DIMX / _DIMY / _DIMZ = nccl | host overrides the staging of that dimension's messages.

# IGG_B200_TIMEOUT_S (600): how long host-side collectives / p2p calls wait for a rank that never arrives.
# Kernel tuning knobs (defaults are the measured best on B200; see profiles/):
#   IGG_TMA_RPT    4  = rows per consumer thread (16 consumer warps + 1 producer warp per CTA)
#   IGG_TILE_BAND  0  = row-major tile order (N > 0: bands of N tile columns)
#   IGG_L2_PROMO   1  = 64-byte L2 promotion for TMA loads (0 none, 2 128 B, 3 256 B)
#   IGG_TMA_STRIP  1  = whole-interior steps of grids with nx >= 512 use the strip-marching kernel (0: tiles)
#   IGG_STRIP_ROWS 256 = rows per work item of the strip-marching kernel (multiple of 32)
#   IGG_RESIDENT_LL 1 = cluster-resident kernel exchanges halos as value+epoch-tag words (0: fence + flag)
#   IGG_E2E_SPIN   1  = step_controlled waits by polling an event (0: cudaStreamSynchronize)
#   IGG_T2_RPT     2  = two-steps-per-pass kernel (experimental): rows per consumer thread (2: 16 warps, 4: 8 warps)
#   IGG_T2_ROWS    252 = its output rows per work item ((rows + 4) % 16 == 0)
#   IGG_T2_SRB     16 = its rows per block (8: half-size rings, two CTAs per SM)
echo "ENV setup done (transport: $IGG_B200_TRANSPORT)"
