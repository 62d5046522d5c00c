#!/bin/bash
cd "$(dirname "${BASH_SOURCE[0]}")/.." || exit 1
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_temporal_gpu.py tests/test_fabric_gpu.py tests/test_kernels_gpu.py -m gpu -x -q 2>&1 | tail -n 3
timeout 300 python scripts/check_temporal.py --skip-timing 2>&1 | tail -n 4
for p in "0 0" "1 0" "0 1" "1 1"; do python scripts/prof_periodic.py $p temporal 40; done 2>&1 | tee $OUT/periodic2.txt
exit 0
