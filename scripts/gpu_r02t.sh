#!/bin/bash
set -u
echo "--- new"; timeout 300 python scripts/sweep_shapes.py c4 2>&1 | grep '"n_ctas": 148'
echo "--- base"; B200ATTN_LIB=sglang_custom_b200/libb200attn_base.so timeout 300 python scripts/sweep_shapes.py c4 2>&1 | grep '"n_ctas": 148'
timeout 600 python -m pytest tests/test_gpu_decode.py -m gpu -x -q 2>&1 | tail -2
