#!/bin/bash
# 1-GPU: the self-fed variant of the two-step kernel (no producer warp): numerics, then A/B against the producer kernel
cd "$(dirname "${BASH_SOURCE[0]}")/.." || exit 1
OUT=gpurun_out; mkdir -p $OUT
for srb in 4 8; do
  IGG_T2_SELF_FED=1 IGG_T2_SRB=$srb timeout 300 python scripts/check_temporal.py --skip-timing > $OUT/sf_check_srb$srb.log 2>&1; echo "self-fed srb=$srb check rc=$?"; tail -n 3 $OUT/sf_check_srb$srb.log
done
IGG_T2_SELF_FED=1 timeout 600 python -m pytest tests/test_temporal_gpu.py tests/test_kernels_gpu.py tests/test_fabric_gpu.py -m gpu -x -q 2>&1 | tail -n 3
for rep in 1 2; do
for sf in 1 0; do
  IGG_T2_SELF_FED=$sf timeout 300 python bench.py --no-baseline --no-verify --steps 50 > $OUT/sf_bench_${sf}_$rep.json 2>/dev/null
  python -c "
import json
d=json.loads([l for l in open('$OUT/sf_bench_${sf}_$rep.json') if l.startswith('{')][-1])
print('rep $rep self_fed=$sf', 'ms/step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['clocks'])"
done
done 2>&1 | tee $OUT/sweep4.txt
exit 0
