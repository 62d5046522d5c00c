#!/bin/bash
# 1-GPU A/B of compile-time / run-time variants of the two-step kernel: CTAs per SM (4 vs 5) x rows per work item
cd "$(dirname "${BASH_SOURCE[0]}")/.." || exit 1
OUT=gpurun_out; mkdir -p $OUT
for rep in 1 2; do
for lib in default ctas5; do
 for rows in 252 380 764; do
  L=""; [ "$lib" != "default" ] && L="$PWD/rocm_mpi_b200/_lib/libigg_b200_$lib.so"
  echo -n "rep=$rep lib=$lib rows=$rows: "
  IGG_B200_LIB=$L IGG_T2_ROWS=$rows timeout 120 python scripts/prof_step.py temporal 16384 16384 62 | tail -n 1
 done
done
done 2>&1 | tee $OUT/sweep2.txt
for lib in default ctas5; do
  L=""; [ "$lib" != "default" ] && L="$PWD/rocm_mpi_b200/_lib/libigg_b200_$lib.so"
  IGG_B200_LIB=$L timeout 300 python bench.py --no-baseline --no-verify --steps 100 > $OUT/sweep2_bench_$lib.json 2>/dev/null
  python -c "
import json
d=json.loads([l for l in open('$OUT/sweep2_bench_$lib.json') if l.startswith('{')][-1])
print('$lib', 'ms/step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['clocks'])"
done
exit 0
