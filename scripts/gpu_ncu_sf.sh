#!/bin/bash
# ncu --set full of the benchmarked two-step kernel (self-fed pipeline), 16384^2, 1 GPU
cd "$(dirname "${BASH_SOURCE[0]}")/.." || exit 1
OUT=gpurun_out; mkdir -p $OUT
python scripts/prof_step.py temporal 16384 16384 6 > $OUT/sf_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:temporal2 -s 2 -c 2 -f -o $OUT/prof_t2_self_fed \
    python scripts/prof_step.py temporal 16384 16384 6 > $OUT/sf_ncu.log 2>&1
cat $OUT/sf_plain.log; tail -n 3 $OUT/sf_ncu.log
exit 0
