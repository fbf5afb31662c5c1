#!/bin/bash
# usage: profiles/run_multi.sh N [quick]   (under gpurun --gpus N): all-reduce parity at N ranks + bench lines at TP=N
# quick: only the all-reduce test and the headline bench
N=$1
QUICK=$2
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$TR --master-port 29541 tests/multi_gpu_allreduce.py > gpurun_out/r02_allreduce_tp$N.txt 2>&1; echo "allreduce rc=$?"; tail -4 gpurun_out/r02_allreduce_tp$N.txt
$TR --master-port 29542 bench.py --gpus $N > gpurun_out/r02_bench_b_tp$N.json 2> gpurun_out/r02_bench_b_tp$N.err; echo "bench b rc=$?"; cat gpurun_out/r02_bench_b_tp$N.json | cut -c1-400
if [ -z "$QUICK" ] && { [ "$N" = "2" ] || [ "$N" = "4" ]; }; then
  $TR --master-port 29543 bench.py --gpus $N --config c --steps 6 --warmup 3 > gpurun_out/r02_bench_c_tp$N.json 2> gpurun_out/r02_bench_c_tp$N.err; echo "bench c rc=$?"; cat gpurun_out/r02_bench_c_tp$N.json | cut -c1-300
fi
if [ -z "$QUICK" ] && [ "$N" = "8" ]; then
  $TR --master-port 29544 bench.py --gpus $N --config e --steps 64 > gpurun_out/r02_bench_e_tp$N.json 2> gpurun_out/r02_bench_e_tp$N.err; echo "bench e rc=$?"; cat gpurun_out/r02_bench_e_tp$N.json | cut -c1-400
fi
