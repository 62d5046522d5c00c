#!/bin/bash
# compute-sanitizer on the single-GPU periodic self-neighbour configuration, where the tool sees every access of the
# push / flag protocol (one tool per gpurun call).   usage: scripts/gpu_sanitizer.sh memcheck|racecheck
cd "$(dirname "${BASH_SOURCE[0]}")/.." || exit 1
TOOL=${1:-memcheck}
OUT=gpurun_out; mkdir -p $OUT
T="tests/test_fabric_gpu.py::test_strip_kernel_periodic_self_neighbor tests/test_temporal_gpu.py::test_temporal_mode_equals_fused_mode_with_self_neighbour"
timeout 300 python -m pytest $T -x -q > $OUT/san_plain_$TOOL.log 2>&1 || { echo "plain run failed"; tail -n 5 $OUT/san_plain_$TOOL.log; exit 1; }
tail -n 2 $OUT/san_plain_$TOOL.log
timeout 1200 compute-sanitizer --tool $TOOL --log-file $OUT/sanitizer_$TOOL.txt python -m pytest $T -x -q > $OUT/san_run_$TOOL.log 2>&1
echo "sanitizer rc=$?"; tail -n 3 $OUT/san_run_$TOOL.log; tail -n 15 $OUT/sanitizer_$TOOL.txt
exit 0
