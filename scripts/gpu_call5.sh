#!/bin/bash
# round-2 GPU call 5 (1 GPU): gpu test tier (kp single-launch paths included), bench with the 4-row two-step kernel, ncu of it
cd "$(dirname "${BASH_SOURCE[0]}")/.." || exit 1
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/c5_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 8 $OUT/c5_pytest.log
timeout 400 python bench.py --steps 50 --warmup 10 > $OUT/c5_bench.json 2> $OUT/c5_bench.err; echo "bench rc=$?"; cat $OUT/c5_bench.json; tail -n 3 $OUT/c5_bench.err
python scripts/prof_step.py temporal 16384 16384 6 > $OUT/c5_plain_t2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:temporal2 -s 2 -c 2 -f -o $OUT/prof_t2_lean_srb4 \
    python scripts/prof_step.py temporal 16384 16384 6 > $OUT/c5_ncu_t2.log 2>&1
cat $OUT/c5_plain_t2.log; tail -n 3 $OUT/c5_ncu_t2.log
exit 0
