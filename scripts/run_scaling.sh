#!/bin/bash
# Weak-scaling run on one node: bench.py at N = 1, 2, 4, 8 GPUs plus the comparison arms at the largest N.
#   scripts/run_scaling.sh [maxN] [steps]      -> gpurun_out/scaling/*.json
set -u
MAXN=${1:-8}; STEPS=${2:-100}
OUT=gpurun_out/scaling; mkdir -p $OUT
run() {  # N impl extra...
  local n=$1 impl=$2; shift 2
  local devs=$(seq -s, 0 $((n-1)))
  if [ "$n" = "1" ]; then
    CUDA_VISIBLE_DEVICES=0 timeout 200 python bench.py --gpus 1 --steps $STEPS --warmup 10 --impl $impl "$@" 2>$OUT/err_${impl}_$n.log | tail -1 > $OUT/bench_${impl}_$n.json
  else
    CUDA_VISIBLE_DEVICES=$devs timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 \
      bench.py --gpus $n --steps $STEPS --warmup 10 --impl $impl "$@" 2>$OUT/err_${impl}_$n.log | tail -1 > $OUT/bench_${impl}_$n.json
  fi
  python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_${impl}_$n.json"))
    e=d.get("e2e") or {}
    print("$impl N=$n ms/step", d["ms_per_step"], "T_eff/GPU", d["per_gpu_T_eff"], "aggregate", d["value"], "e2e ms", e.get("ms_per_step"), "launches", d["gpu_launches"], "clocks", d.get("clocks"))
except Exception as ex:
    print("$impl N=$n FAILED", ex)
PY
}
for n in 1 2 4 8; do [ $n -le $MAXN ] && run $n ours; done
for impl in ${ARMS:-nccl nccl_tma plain_peer hide}; do run $MAXN $impl --no-e2e; done
