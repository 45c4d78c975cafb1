#!/bin/bash
set -u
for lv in 1 2; do echo "--- B200ATTN_DECODE_FP8_BIG=$lv"; B200ATTN_DECODE_FP8_BIG=$lv timeout 300 python scripts/sweep_shapes.py lowbatch 2>&1 | grep "fp8" | grep '"n_ctas": 148'; B200ATTN_DECODE_FP8_BIG=$lv timeout 300 python scripts/sweep_shapes.py c4 2>&1 | grep "tp8" | grep '"n_ctas": 148'; done
