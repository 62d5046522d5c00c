#!/bin/bash
# 1-GPU: temporal tests, then A/B of 5 vs 4 CTAs per SM (device-only and end-to-end with the in-kernel diagnostics)
cd "$(dirname "${BASH_SOURCE[0]}")/.." || exit 1
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_temporal_gpu.py tests/test_kernels_gpu.py -m gpu -x -q 2>&1 | tail -n 3
for rep in 1 2; do
for lib in default ctas4; do
  L=""; [ "$lib" != "default" ] && L="$PWD/rocm_mpi_b200/_lib/libigg_b200_$lib.so"
  IGG_B200_LIB=$L timeout 300 python bench.py --no-baseline --no-verify --steps 50 > $OUT/sweep3_bench_${lib}_$rep.json 2>/dev/null
  python -c "
import json
d=json.loads([l for l in open('$OUT/sweep3_bench_${lib}_$rep.json') if l.startswith('{')][-1])
print('rep $rep', '$lib', 'ms/step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['clocks'])"
done
done 2>&1 | tee $OUT/sweep3.txt
exit 0
