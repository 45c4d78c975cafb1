#!/bin/bash
# One gpurun call: GPU parity tests, bench line, ncu launch list, ncu --set full of the two hot kernels.
#   gpurun --timeout 1500 -- 'bash scripts/gpu_round.sh'
set -u
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
timeout 600 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
cat $O/bench.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?"
cat $O/bench_ref.json
# launch list of the bench command (cold-cache, serialised): shares only
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file $O/launches_bench.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu > $O/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
# full captures
timeout 400 ncu --set full --clock-control none --import-source on -k regex:decode_kernel -s 2 -c 1 -f -o $O/decode_c2 \
    python scripts/prof_kernels.py decode 4 > $O/ncu_decode.log 2>&1; echo "ncu decode rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:extend_tc -s 2 -c 1 -f -o $O/extend_c3 \
    python scripts/prof_kernels.py extend 4 > $O/ncu_extend.log 2>&1; echo "ncu extend rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:decode_kernel -s 2 -c 1 -f -o $O/decode_c4_fp8 \
    python scripts/prof_kernels.py decode_fp8 4 > $O/ncu_decode_fp8.log 2>&1; echo "ncu decode fp8 rc=$?"
ls -la $O
