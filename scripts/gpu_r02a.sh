#!/bin/bash
# round 2, call A: GPU parity tests, the new bench line at N = 1, same-box GPU baselines
set -u
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_r02a.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_gpu_r02a.log
tail -5 $O/pytest_gpu_r02a.log
timeout 900 python bench.py > $O/bench_r02a.json 2> $O/bench_r02a.err; echo "bench rc=$?"
tail -c 3000 $O/bench_r02a.json
tail -5 $O/bench_r02a.err
timeout 1500 python scripts/time_gpu_baselines.py --timeout 300 --out $O/gpu_baselines_r02a.jsonl > $O/gpu_baselines_r02a.log 2>&1; echo "baselines rc=$?"
cat $O/gpu_baselines_r02a.jsonl
