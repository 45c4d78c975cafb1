#!/bin/bash
# round 2, call B: token-major FP8 decode consumer -- parity, then A/B timing at the C4 per-rank shapes
set -u
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_decode.py tests/test_gpu_masked.py tests/test_gpu_tp.py tests/test_gpu_extend.py -m gpu -x -q > $O/pytest_gpu_r02b.log 2>&1; echo "pytest rc=$?"
tail -5 $O/pytest_gpu_r02b.log
echo "--- token-major (default)"; timeout 600 python scripts/sweep_shapes.py c4 2>&1 | tee $O/sweep_c4_tokm1.txt
echo "--- head-major (old)"; B200ATTN_DECODE_FP8_TOKM=0 timeout 600 python scripts/sweep_shapes.py c4 2>&1 | tee $O/sweep_c4_tokm0.txt
echo "--- layout"; timeout 600 python scripts/sweep_shapes.py c4layout 2>&1 | tee $O/sweep_c4layout_tokm1.txt
