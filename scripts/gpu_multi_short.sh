#!/bin/bash
# short multi-GPU check of a change to the two-step path: numerics on real neighbours, then the bench line
# (IGG_T2_LEAD=0/1 A/B), optionally the gather benchmark.   usage: scripts/gpu_multi_short.sh N [gather]
cd "$(dirname "${BASH_SOURCE[0]}")/.." || exit 1
N=${1:-2}
DIMS=${DIMS:-}                      # e.g. DIMS="1 2": neighbours in y only (strided faces)
CT_DIMS=""; BE_DIMS=""
if [ -n "$DIMS" ]; then set -- $DIMS; CT_DIMS="--dimx $1 --dimy $2"; BE_DIMS="--dims $1 $2"; set -- $N; fi
OUT=gpurun_out; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541"
timeout 300 $TR scripts/check_temporal.py --skip-timing $CT_DIMS > $OUT/s${N}_check_t2.log 2>&1; echo "check rc=$?"; grep -E "steps|OK|MISMATCH|rror" $OUT/s${N}_check_t2.log | tail -n 5
for lead in 1 0; do
  IGG_T2_LEAD=$lead timeout 300 $TR bench.py --gpus $N --steps 100 --warmup 10 --no-baseline --no-verify $BE_DIMS > $OUT/s${N}_bench_lead$lead.json 2> $OUT/s${N}_bench_lead$lead.err
  echo "lead=$lead rc=$?"; python -c "
import json,sys
d=json.loads([l for l in open('$OUT/s${N}_bench_lead$lead.json') if l.startswith('{')][-1])
print('  ms/step', d['ms_per_step'], 'aggregate', d['value'], 'e2e ms', d['e2e']['ms_per_step'], 'clocks', d['clocks'])"
done
timeout 400 $TR bench.py --gpus $N --steps 50 --warmup 10 $BE_DIMS > $OUT/s${N}_bench_full.json 2> $OUT/s${N}_bench_full.err; echo "full bench rc=$?"; cut -c1-330 $OUT/s${N}_bench_full.json
if [ -n "${2:-}" ]; then
  timeout 300 $TR benchmarks/gather_bench.py --size 12288 --out $OUT/s${N}_gather.json > $OUT/s${N}_gather.log 2>&1; echo "gather rc=$?"; grep "^{" $OUT/s${N}_gather.log | tail -n 1
fi
exit 0
