#!/bin/bash
# ncu --set full of the benchmarked two-step kernel (self-fed pipeline), 16384^2, 1 GPU
cd "$(dirname "${BASH_SOURCE[0]}")/.." || exit 1
OUT=gpurun_out; mkdir -p $OUT
python scripts/prof_step.py temporal 16384 16384 6 > $OUT/sf_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:temporal2 -s 2 -c 2 -f -o $OUT/prof_t2_self_fed \
    python scripts/prof_step.py temporal 16384 16384 6 > $OUT/sf_ncu.log 2>&1
cat $OUT/sf_plain.log; tail -n 3 $OUT/sf_ncu.log
IGG_T2_SRB=16 timeout 300 python scripts/check_temporal.py --skip-timing > $OUT/sf_check_srb16.log 2>&1; echo "self-fed srb=16 check rc=$?"; tail -n 2 $OUT/sf_check_srb16.log
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 3
exit 0
