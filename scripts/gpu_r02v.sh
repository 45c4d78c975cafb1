#!/bin/bash
set -u
echo "--- level 2 (HB >= 2)"; B200ATTN_DECODE_FP8_BIG=2 timeout 300 python scripts/sweep_shapes.py c4 2>&1 | grep '"n_ctas": 148'
B200ATTN_DECODE_FP8_BIG=2 timeout 600 python -m pytest tests/test_gpu_decode.py tests/test_gpu_tp.py -m gpu -x -q 2>&1 | tail -2
echo "--- low batch, level 1 vs 0"
timeout 300 python scripts/sweep_shapes.py lowbatch 2>&1 | tail -12
B200ATTN_DECODE_FP8_BIG=0 timeout 300 python scripts/sweep_shapes.py lowbatch 2>&1 | tail -12
