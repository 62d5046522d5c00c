#!/bin/bash
# round-2 multi-GPU evidence run on N GPUs of one box (2, 4 or 8): two-step kernel numerics on real neighbours,
# the multigpu test tier, bench.py (verify + same-box baselines), the update_halo API sweep, the contiguous p2p
# sweep, small-grid latency incl. the kp paths, the gather benchmark, and (N = 8) a 3-D weak-scaling line.
#   usage: scripts/gpu_multi.sh N [quick]        -> gpurun_out/m<N>_*
cd "$(dirname "${BASH_SOURCE[0]}")/.." || exit 1
N=${1:-2}; QUICK=${2:-}
OUT=gpurun_out; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541"
nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm --format=csv,noheader | head -n $N
free -g | head -n 2
echo "== check_temporal on $N ranks"
timeout 300 $TR scripts/check_temporal.py --skip-timing > $OUT/m${N}_check_t2.log 2>&1; echo "rc=$?"; grep -E "steps|OK|MISMATCH|rror" $OUT/m${N}_check_t2.log | tail -n 6
if [ -z "$QUICK" ]; then
  echo "== multigpu tests"
  timeout 900 python -m pytest tests/test_multi_gpu.py -m multigpu -x -q > $OUT/m${N}_pytest_mg.log 2>&1; echo "rc=$?"; tail -n 4 $OUT/m${N}_pytest_mg.log
fi
echo "== bench ours (two-step) N=$N"
timeout 400 $TR bench.py --gpus $N --steps 50 --warmup 10 > $OUT/m${N}_bench_ours.json 2> $OUT/m${N}_bench_ours.err; echo "rc=$?"; cut -c1-330 $OUT/m${N}_bench_ours.json; grep -v "OMP_NUM\|\*\*\*\*" $OUT/m${N}_bench_ours.err | tail -n 3
echo "== bench fused (one step per pass) N=$N"
timeout 400 $TR bench.py --gpus $N --steps 50 --warmup 10 --kernel fused --no-baseline > $OUT/m${N}_bench_fused.json 2> $OUT/m${N}_bench_fused.err; echo "rc=$?"; cut -c1-330 $OUT/m${N}_bench_fused.json
[ -z "$QUICK" ] && echo "== bench hide (frame || interior on two priority streams, b_width 16 4) N=$N" && \
timeout 400 $TR bench.py --gpus $N --steps 50 --warmup 10 --impl hide --no-baseline > $OUT/m${N}_bench_hide.json 2> $OUT/m${N}_bench_hide.err; echo "rc=$?"; cut -c1-330 $OUT/m${N}_bench_hide.json
echo "== update_halo API sweep N=$N"
timeout 400 $TR benchmarks/update_halo_sweep.py --out $OUT/m${N}_update_halo_sweep.json 2>&1 | grep "^{" | tail -n 8
echo "== p2p sweep N=$N"
timeout 300 $TR benchmarks/halo_bw_sweep.py --max-mb 256 --out $OUT/m${N}_halo_bw.json 2>&1 | grep "^{" | tail -n 10
echo "== small-grid latency (global 254 x 254) N=$N"
case $N in 2) L="--nx 128 --ny 254 --dimx 2 --dimy 1";; 4) L="--nx 128 --ny 128 --dimx 2 --dimy 2";; *) L="--nx 65 --ny 128 --dimx 4 --dimy 2";; esac
timeout 300 $TR benchmarks/small_grid_latency.py $L --out $OUT/m${N}_small_254.json 2>&1 | grep "^{" | tail -n 1
echo "== small-grid latency, literal 252 x 252 local N=$N"
timeout 300 $TR benchmarks/small_grid_latency.py --nx 252 --ny 252 --out $OUT/m${N}_small_252.json 2>&1 | grep "^{" | tail -n 1
echo "== gather bench N=$N"
timeout 300 $TR benchmarks/gather_bench.py --size 12288 --out $OUT/m${N}_gather.json > $OUT/m${N}_gather.log 2>&1; echo "rc=$?"; grep "^{" $OUT/m${N}_gather.log | tail -n 1; grep -E "Error|error|Traceback" -A3 $OUT/m${N}_gather.log | grep -v Warning | tail -n 12
if [ "$N" = "8" ]; then
  echo "== 3-D weak scaling 512^3 per GPU on 2x2x2"
  timeout 300 $TR -m rocm_mpi_b200.models.diffusion3d --nx 512 --ny 512 --nz 512 --nt 60 --warmup 10 --dimx 2 --dimy 2 --dimz 2 --do-vis 0 --quiet 0 2>&1 | grep -E "Executed|Global grid" | tail -n 2 | tee $OUT/m${N}_3d_512.txt
  CUDA_VISIBLE_DEVICES=0 timeout 200 python -m rocm_mpi_b200.models.diffusion3d --nx 512 --ny 512 --nz 512 --nt 60 --warmup 10 --do-vis 0 --quiet 0 2>&1 | grep -E "Executed|Global grid" | tail -n 2 | tee $OUT/m1_3d_512.txt
fi
exit 0
