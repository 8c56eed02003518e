#!/bin/bash
# Multi-GPU visit (under `gpurun --gpus N`): headline bench and the C4 training step at N ranks, one process per GPU.
# Usage: bash scripts/gpu_scaling.sh <tag> <N> [extra bench.py flags for the headline run]
TAG=${1:-rXX}; N=${2:-2}; shift 2
O=gpurun_out; mkdir -p $O
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 600 $RUN bench.py --gpus $N --steps 5 --warmup 3 "$@" > $O/${TAG}_bench_n$N.json 2> $O/${TAG}_bench_n$N.err; echo "bench n=$N rc=$?"
timeout 400 $RUN bench.py --workload c4 --gpus $N --steps 5 --warmup 3 > $O/${TAG}_train_bench_c4_n$N.json 2> $O/${TAG}_train_n$N.err; echo "train n=$N rc=$?"
tail -c 600 $O/${TAG}_train_bench_c4_n$N.json
