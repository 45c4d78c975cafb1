#!/bin/bash
# round 2, call W: ncu --set full of the 64 KiB-stage FP8 decode kernel at C4 / TP 1 (after), and of the 32 KiB-stage form (before)
set -u
O=gpurun_out
mkdir -p $O
timeout 400 ncu --set full --clock-control none --import-source on -k regex:decode_kernel -s 2 -c 1 -f -o $O/decode_c4_fp8_tp1_big python scripts/prof_kernels.py decode_fp8_tp1 4 > $O/ncu_fp8_tp1_big.log 2>&1; echo "ncu big rc=$?"
B200ATTN_DECODE_FP8_BIG=0 timeout 400 ncu --set full --clock-control none --import-source on -k regex:decode_kernel -s 2 -c 1 -f -o $O/decode_c4_fp8_tp1_small python scripts/prof_kernels.py decode_fp8_tp1 4 > $O/ncu_fp8_tp1_small.log 2>&1; echo "ncu small rc=$?"
