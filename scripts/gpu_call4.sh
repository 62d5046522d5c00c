#!/bin/bash
# round-2 multi-GPU call (N GPUs, default 2): multi-rank numerics of the two-step kernel, multigpu test tier, bench with
# verify + baselines, update_halo API sweep, contiguous p2p sweep.   usage: gpu_call4.sh [N]
cd "$(dirname "${BASH_SOURCE[0]}")/.." || exit 1
N=${1:-2}
OUT=gpurun_out; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541"
echo "== check_temporal on $N ranks"
timeout 300 $TR scripts/check_temporal.py --skip-timing > $OUT/c4_check_t2_$N.log 2>&1; echo "rc=$?"; grep -E "steps|OK|MISMATCH|rror" $OUT/c4_check_t2_$N.log | tail -n 8
echo "== multigpu tests"
timeout 900 python -m pytest tests/test_multi_gpu.py -m multigpu -x -q > $OUT/c4_pytest_mg_$N.log 2>&1; echo "rc=$?"; tail -n 4 $OUT/c4_pytest_mg_$N.log
echo "== bench ours (two-step) N=$N"
timeout 400 $TR bench.py --gpus $N --steps 50 --warmup 10 > $OUT/c4_bench_ours_$N.json 2> $OUT/c4_bench_ours_$N.err; echo "rc=$?"; cat $OUT/c4_bench_ours_$N.json; tail -n 3 $OUT/c4_bench_ours_$N.err
echo "== bench fused (one step per pass) N=$N"
timeout 400 $TR bench.py --gpus $N --steps 50 --warmup 10 --kernel fused --no-baseline > $OUT/c4_bench_fused_$N.json 2> $OUT/c4_bench_fused_$N.err; echo "rc=$?"; cat $OUT/c4_bench_fused_$N.json
echo "== update_halo API sweep N=$N"
timeout 400 $TR benchmarks/update_halo_sweep.py --out $OUT/c4_update_halo_sweep_$N.json 2>&1 | grep -v Warning | tail -n 8
echo "== p2p sweep N=$N"
timeout 300 $TR benchmarks/halo_bw_sweep.py --max-mb 256 --out $OUT/c4_halo_bw_$N.json 2>&1 | grep -v Warning | tail -n 10
exit 0
