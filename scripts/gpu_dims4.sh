#!/bin/bash
# 4 GPUs: where does the two-step kernel lose time with neighbours -- x faces (contiguous rows) or y faces (strided columns)?
cd "$(dirname "${BASH_SOURCE[0]}")/.." || exit 1
OUT=gpurun_out; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29541"
for d in "4 1" "1 4" "2 2"; do
  for kern in temporal fused; do
  timeout 300 $TR bench.py --gpus 4 --steps 50 --warmup 10 --no-baseline --no-verify --no-e2e --kernel $kern --dims $d > $OUT/d4.json 2>/dev/null
  python -c "
import json
d=json.loads([l for l in open('$OUT/d4.json') if l.startswith('{')][-1])
print('dims $d $kern', 'ms/step', d['ms_per_step'], d['config']['parallelism'], d['clocks'])"
  done
done 2>&1 | tee $OUT/dims4.txt
exit 0
