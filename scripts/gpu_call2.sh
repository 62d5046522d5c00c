#!/bin/bash
# round-2 GPU call 2 (1 GPU): gpu test tier, new bench.py (verify + baselines + pipelined e2e), ncu of the two-step kernel
cd "$(dirname "${BASH_SOURCE[0]}")/.." || exit 1
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/c2_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 6 $OUT/c2_pytest.log
timeout 400 python bench.py --steps 50 --warmup 10 > $OUT/c2_bench_temporal.json 2> $OUT/c2_bench_temporal.err; echo "bench rc=$?"; cat $OUT/c2_bench_temporal.json; tail -n 5 $OUT/c2_bench_temporal.err
timeout 400 python bench.py --kernel fused --steps 50 --warmup 10 --no-verify --no-baseline > $OUT/c2_bench_fused.json 2> $OUT/c2_bench_fused.err; echo "bench fused rc=$?"; cat $OUT/c2_bench_fused.json
python scripts/prof_step.py temporal 16384 16384 6 > $OUT/c2_plain_t2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:temporal2 -s 2 -c 2 -f -o $OUT/prof_t2_srb8 \
    python scripts/prof_step.py temporal 16384 16384 6 > $OUT/c2_ncu_t2.log 2>&1
cat $OUT/c2_plain_t2.log; tail -n 3 $OUT/c2_ncu_t2.log
exit 0
