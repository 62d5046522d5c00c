#!/bin/bash
# round-2 GPU call 7 (1 GPU): full gpu tier, gather bench (copy kernel locality), smoke(), compute-sanitizer memcheck
cd "$(dirname "${BASH_SOURCE[0]}")/.." || exit 1
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q > $OUT/c7_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 12 $OUT/c7_pytest.log
timeout 200 python __graft_entry__.py smoke 2>&1 | tail -n 4
timeout 300 python benchmarks/gather_bench.py --size 12288 --out $OUT/c7_gather_1.json 2>&1 | tail -n 1
timeout 300 python benchmarks/halo_bw_sweep.py --max-mb 64 2>&1 | grep "^{" | head -n 3
bash scripts/gpu_sanitizer.sh memcheck
exit 0
