#!/bin/bash
# One GPU visit: parity tests, smoke, bench (both arms), stage timers, ncu launch list, ncu full capture of the decoder kernel.
# Usage (under gpurun): bash scripts/gpu_round.sh <tag> [kernel-regex]
TAG=${1:-rXX}
KREGEX=${2:-decoder3_kernel}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/${TAG}_smi.txt 2>&1
nproc >> $O/${TAG}_smi.txt
timeout 1200 python -m pytest tests -m gpu -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/${TAG}_pytest.log; tail -3 $O/${TAG}_pytest.log
timeout 200 python __graft_entry__.py smoke > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/${TAG}_smoke.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 --ref-budget-s 40 > $O/${TAG}_bench_reference.json 2> $O/${TAG}_bench_reference.err; echo "bench reference rc=$?"
timeout 900 python bench.py --steps 5 --warmup 3 > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"
tail -c 1500 $O/${TAG}_bench.json
timeout 300 python bench.py --workload c4 --steps 3 --warmup 2 > $O/${TAG}_train_bench_c4.json 2> $O/${TAG}_train.err; echo "train bench rc=$?"
TACO_DEC3_DBG=513 timeout 200 python scripts/dec3_timing.py 64 200 > $O/${TAG}_dec3_timing_nowait.txt 2>&1
timeout 200 python scripts/dec3_timing.py 64 200 > $O/${TAG}_dec3_timing.txt 2>&1; head -3 $O/${TAG}_dec3_timing.txt | cut -c1-400
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extra > $O/${TAG}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:${KREGEX} -s 1 -c 1 -o $O/${TAG}_decoder_full -f \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extra > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la $O | tail -20
