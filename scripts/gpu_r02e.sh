#!/bin/bash
# round 2, call E (2 GPUs): TP parity test of the CUDA path at world 2 + the strong-scaling bench line at N = 2
set -u
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi_r02e.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_tp.py -m gpu -x -q > $O/pytest_tp_r02e.log 2>&1; echo "pytest tp rc=$?"
tail -5 $O/pytest_tp_r02e.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 10 --warmup 3 > $O/bench_r02e_n2.json 2> $O/bench_r02e_n2.err; echo "bench n2 rc=$?"
tail -c 2500 $O/bench_r02e_n2.json
grep -v "^\[W" $O/bench_r02e_n2.err | tail -5
