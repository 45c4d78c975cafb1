#!/bin/bash
# round 2: strong-scaling bench line at N GPUs (argument), final library
set -u
N=${1:-2}
O=gpurun_out
mkdir -p $O
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 10 --warmup 3 > $O/bench_final_n$N.json 2> $O/bench_final_n$N.err; echo "bench n$N rc=$?"
tail -c 300 $O/bench_final_n$N.json
