#!/bin/bash
set -u
echo "--- big tiles"; timeout 300 python scripts/sweep_shapes.py c4 2>&1 | grep '"n_ctas": 148'
echo "--- two tile groups"; B200ATTN_DECODE_FP8_BIG=0 timeout 300 python scripts/sweep_shapes.py c4 2>&1 | grep '"n_ctas": 148'
timeout 600 python -m pytest tests/test_gpu_decode.py tests/test_gpu_draft.py tests/test_gpu_masked.py -m gpu -x -q 2>&1 | tail -2
