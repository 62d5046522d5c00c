#!/bin/bash
# round-2 GPU call 3 (1 GPU): lean inner path of the two-step kernel -- numerics, then timing sweep
cd "$(dirname "${BASH_SOURCE[0]}")/.." || exit 1
OUT=gpurun_out; mkdir -p $OUT
for srb in 8 4 16; do
  echo "== numerics IGG_T2_SRB=$srb"
  IGG_T2_SRB=$srb timeout 300 python scripts/check_temporal.py --skip-timing > $OUT/c3_check_srb$srb.log 2>&1; echo "rc=$?"; tail -n 4 $OUT/c3_check_srb$srb.log
done
timeout 300 python -m pytest tests/test_temporal_gpu.py -m gpu -x -q 2>&1 | tail -n 3
for srb in 4 8 16; do for lean in 1 0; do for rows in 252 508; do
  echo "== IGG_T2_SRB=$srb IGG_T2_LEAN=$lean IGG_T2_ROWS=$rows"
  IGG_T2_SRB=$srb IGG_T2_LEAN=$lean IGG_T2_ROWS=$rows timeout 120 python scripts/prof_step.py temporal 16384 16384 42
done; done; done 2>&1 | tee $OUT/c3_sweep.txt
timeout 400 python bench.py --steps 50 --warmup 10 --no-baseline > $OUT/c3_bench.json 2> $OUT/c3_bench.err; echo "bench rc=$?"; cat $OUT/c3_bench.json
exit 0
