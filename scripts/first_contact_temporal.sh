#!/bin/bash
# First GPU contact of the staged two-steps-per-pass kernel, as ONE gpurun command (1 GPU):
#   /usr/local/graft/bin/gpurun --timeout 2400 -- 'bash scripts/first_contact_temporal.sh'
# Every stage runs under its own timeout and writes to gpurun_out/; a failing stage does not stop the later ones,
# except that timing and the sweep are skipped when the numerics check failed.
REPO="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
cd "$REPO" || exit 1
OUT=gpurun_out
mkdir -p $OUT
echo "== stage 1: numerics (kernel alone, solver with self-neighbour, controlled pairs)" | tee $OUT/t2_summary.txt
timeout 300 python scripts/check_temporal.py --skip-timing > $OUT/t2_check.log 2>&1
rc=$?; echo "check_temporal rc=$rc" | tee -a $OUT/t2_summary.txt; tail -n 25 $OUT/t2_check.log
echo "== stage 2: staged pytest tier" | tee -a $OUT/t2_summary.txt
timeout 400 python -m pytest tests/test_temporal_gpu.py tests/test_hide_drivers_gpu.py -m gpu -q > $OUT/t2_pytest.log 2>&1
echo "pytest rc=$?" | tee -a $OUT/t2_summary.txt; tail -n 5 $OUT/t2_pytest.log | tee -a $OUT/t2_summary.txt
if [ $rc -eq 0 ]; then
  echo "== stage 3: timing at 16384^2, fused vs temporal" | tee -a $OUT/t2_summary.txt
  timeout 300 python scripts/check_temporal.py --timing-only 2>&1 | tee -a $OUT/t2_summary.txt
  echo "== stage 4: sweep" | tee -a $OUT/t2_summary.txt
  timeout 600 bash scripts/sweep_temporal.sh 16384 16384 30 >> $OUT/t2_summary.txt 2>&1
  echo "== stage 5: bench line (value + pair-controlled e2e)" | tee -a $OUT/t2_summary.txt
  timeout 300 python bench.py --impl temporal --steps 50 --warmup 10 2>&1 | tail -n 1 | tee $OUT/bench_temporal_1gpu.json
fi
echo "== stage 6: plain-C GPU example" | tee -a $OUT/t2_summary.txt
gcc -std=c99 -O2 -I include -I /usr/local/cuda/include examples/c/heat_gpu.c -o /tmp/heat_gpu \
    -L rocm_mpi_b200/_lib -ligg_b200 -Wl,-rpath,$REPO/rocm_mpi_b200/_lib -L /usr/local/cuda/lib64 -lcudart \
  && timeout 120 /tmp/heat_gpu 2>&1 | tee -a $OUT/t2_summary.txt
exit 0
