#!/bin/bash
# Tuning sweep of the two-steps-per-pass kernel on ONE GPU (the knobs are read once per process, hence one
# process per point).  Run only after scripts/check_temporal.py has printed OK.
#   scripts/sweep_temporal.sh [nx ny steps]  ->  gpurun_out/temporal_sweep.txt
REPO="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
NX=${1:-16384}; NY=${2:-16384}; STEPS=${3:-40}
OUT="$REPO/gpurun_out/temporal_sweep.txt"
mkdir -p "$REPO/gpurun_out"
: > "$OUT"
for srb in 16 8; do
 for rpt in 2 4; do
  for rows in 124 252 508; do
    echo "== IGG_T2_SRB=$srb IGG_T2_RPT=$rpt IGG_T2_ROWS=$rows" | tee -a "$OUT"
    IGG_T2_SRB=$srb IGG_T2_RPT=$rpt IGG_T2_ROWS=$rows timeout 120 python "$REPO/scripts/check_temporal.py" \
        --nx "$NX" --ny "$NY" --steps "$STEPS" --timing-only 2>&1 | grep -E "ms/step|MISMATCH|rror" | tee -a "$OUT"
  done
 done
done
