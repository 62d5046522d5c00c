#!/bin/bash
# Profile the top kernel with Nsight Compute on ONE GPU (run under gpurun).
#   scripts/profile_kernel.sh <tma|direct> <tag>
# 1) plain run must exit 0, 2) launch list with device times, 3) one --set full capture (3 launches).
set -u
variant=${1:-tma}; tag=${2:-$variant}
mkdir -p gpurun_out
python scripts/prof_step.py $variant 16384 16384 6 > gpurun_out/plain_$tag.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_$tag.log; exit 1; }
cat gpurun_out/plain_$tag.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_$tag.csv \
    python scripts/prof_step.py $variant 16384 16384 6 > gpurun_out/ncu_launch_$tag.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:diff2d -s 2 -c 3 -f -o gpurun_out/prof_$tag \
    python scripts/prof_step.py $variant 16384 16384 6 > gpurun_out/ncu_full_$tag.log 2>&1
tail -3 gpurun_out/ncu_full_$tag.log
ls -la gpurun_out/ | head -30
