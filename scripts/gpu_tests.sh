#!/bin/bash
# the driver's round-end GPU checks on one GPU: gpu test tier, smoke(), default bench line
cd "$(dirname "${BASH_SOURCE[0]}")/.." || exit 1
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/t_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 6 $OUT/t_pytest.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 3
timeout 400 python bench.py > $OUT/t_bench.json 2> $OUT/t_bench.err; echo "bench rc=$?"; cut -c1-300 $OUT/t_bench.json
timeout 60 python bench.py --impl reference
exit 0
