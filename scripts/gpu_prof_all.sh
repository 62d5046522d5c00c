#!/bin/bash
# one ncu --set full capture of the FIRST launch of every kernel of the library (1 GPU) -> gpurun_out/prof_all_r2.ncu-rep
cd "$(dirname "${BASH_SOURCE[0]}")/.." || exit 1
OUT=gpurun_out; mkdir -p $OUT
python scripts/prof_all_kernels.py > $OUT/prof_all_plain.log 2>&1 && \
ncu --set full --clock-control none --kernel-id :::1 -f -o $OUT/prof_all_r2 python scripts/prof_all_kernels.py > $OUT/prof_all_ncu.log 2>&1
tail -n 2 $OUT/prof_all_plain.log; tail -n 2 $OUT/prof_all_ncu.log
ncu -i $OUT/prof_all_r2.ncu-rep --page raw --csv > $OUT/prof_all_r2_raw.csv 2>/dev/null
python scripts/summarise_ncu_raw.py $OUT/prof_all_r2_raw.csv | tee $OUT/prof_all_r2_table.md | head -n 40
rm -f $OUT/prof_all_r2.ncu-rep
exit 0
