#!/bin/bash
# Multi-GPU bench lines of round 2 (run under `gpurun --gpus N`): bash profiles/run_r02_multi.sh <N> <tag>
N=${1:-2}
tag=${2:-s38}
out=gpurun_out
mkdir -p $out
run() { name=$1; shift; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 500)) bench.py --gpus $N "$@" > $out/r02_bench_${N}gpu_${name}_$tag.json 2> $out/err_${N}gpu_$name.log; grep -o '"ms_per_step": [0-9.]*' $out/r02_bench_${N}gpu_${name}_$tag.json | head -1; }
if [ "$N" = 2 ]; then XNN_TEST_WATCHDOG=200 timeout 300 python -m pytest tests/test_gpu_nccl.py -x -q -m gpu 2>&1 | tail -2; fi
run default --steps 200 --warmup 5 --no-aux
run syncbn --steps 200 --warmup 5 --no-aux --sync-bn
run strong8192 --steps 200 --warmup 5 --no-aux --global-batch 8192
if [ "$N" = 8 ]; then run cfg4 --workload cfg4 --steps 5 --warmup 3 --no-aux; run cfg5 --workload cfg5 --steps 5 --warmup 3 --no-aux; fi
