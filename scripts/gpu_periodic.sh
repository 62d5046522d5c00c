#!/bin/bash
# 1 GPU: cost of x faces vs y faces of the two-step path with a self-neighbour; per-kernel device times for the y case
cd "$(dirname "${BASH_SOURCE[0]}")/.." || exit 1
OUT=gpurun_out; mkdir -p $OUT
for p in "0 0" "1 0" "0 1" "1 1"; do python scripts/prof_periodic.py $p temporal 40; done 2>&1 | tee $OUT/periodic.txt
for p in "0 0" "0 1"; do python scripts/prof_periodic.py $p fused 40; done 2>&1 | tee -a $OUT/periodic.txt
python scripts/prof_periodic.py 0 1 temporal 6 > $OUT/periodic_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $OUT/periodic_launches.csv \
    python scripts/prof_periodic.py 0 1 temporal 6 > $OUT/periodic_ncu.log 2>&1
grep -E "midlines|temporal2|strip|neighbor|halo_exchange|wait_epoch" $OUT/periodic_launches.csv | awk -F'","' '{print $5, $NF}' | tail -n 16
exit 0
