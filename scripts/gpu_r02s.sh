#!/bin/bash
# round 2, call S (8 GPUs): TP parity test at world 2 / 4 / 8 + the strong-scaling bench line at N = 8
set -u
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_tp.py -m gpu -x -q > $O/pytest_tp_r02s.log 2>&1; echo "pytest tp rc=$?"
tail -3 $O/pytest_tp_r02s.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus 8 --steps 10 --warmup 3 > $O/bench_r02s_n8.json 2> $O/bench_r02s_n8.err; echo "bench n8 rc=$?"
tail -c 1500 $O/bench_r02s_n8.json
grep -v "^\[W\|UserWarning\|return func" $O/bench_r02s_n8.err | tail -5
