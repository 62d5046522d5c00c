#!/bin/bash
# round-2 GPU call 6 (1 GPU): gpu tier with the device gather / kp paths, bench (e2e with the cheaper in-kernel
# diagnostic), small-grid latency rows incl. kp, gather bench on one rank
cd "$(dirname "${BASH_SOURCE[0]}")/.." || exit 1
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/c6_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 8 $OUT/c6_pytest.log
timeout 400 python bench.py --steps 50 --warmup 10 --no-baseline > $OUT/c6_bench.json 2> $OUT/c6_bench.err; echo "bench rc=$?"; cut -c1-400 $OUT/c6_bench.json; tail -n 3 $OUT/c6_bench.err
timeout 300 python benchmarks/small_grid_latency.py --nx 128 --ny 128 --out $OUT/c6_small_128.json 2>&1 | tail -n 2
timeout 300 python benchmarks/small_grid_latency.py --nx 252 --ny 252 --out $OUT/c6_small_252.json 2>&1 | tail -n 2
timeout 300 python benchmarks/gather_bench.py --size 12288 --out $OUT/c6_gather_1.json 2>&1 | tail -n 2
exit 0
