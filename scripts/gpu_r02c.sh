#!/bin/bash
# round 2, call C: token-major FP8 decode + TMA paged-tile loads -- parity first (short time limits), then A/B timing
set -u
O=gpurun_out
mkdir -p $O
timeout 400 python -m pytest tests/test_gpu_decode.py -m gpu -x -q > $O/pytest_decode_r02c.log 2>&1; echo "pytest decode rc=$?"
tail -15 $O/pytest_decode_r02c.log
timeout 400 python -m pytest tests/test_gpu_extend.py -m gpu -x -q > $O/pytest_extend_r02c.log 2>&1; echo "pytest extend rc=$?"
tail -15 $O/pytest_extend_r02c.log
echo "--- token-major (default)"; timeout 200 python scripts/sweep_shapes.py c4 2>&1 | tee $O/sweep_c4_tokm1.txt
echo "--- head-major (old)"; B200ATTN_DECODE_FP8_TOKM=0 timeout 200 python scripts/sweep_shapes.py c4 2>&1 | tee $O/sweep_c4_tokm0.txt
echo "--- layout"; timeout 200 python scripts/sweep_shapes.py c4layout 2>&1 | tee $O/sweep_c4layout_tokm1.txt
timeout 600 python -m pytest tests -m gpu -q > $O/pytest_gpu_r02c.log 2>&1; echo "pytest all rc=$?"
tail -8 $O/pytest_gpu_r02c.log
