#!/bin/bash
# round 2, call D: draft-mode parity + ncu captures of the token-major FP8 decode kernel
set -u
O=gpurun_out
mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_draft.py tests/test_gpu_masked.py -m gpu -x -q > $O/pytest_draft_r02d.log 2>&1; echo "pytest draft rc=$?"
tail -25 $O/pytest_draft_r02d.log
for w in decode_fp8 decode_fp8_arange decode_fp8_tp1; do
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:decode_kernel -s 2 -c 1 -f -o $O/r02_$w python scripts/prof_kernels.py $w 4 > $O/ncu_$w.log 2>&1; echo "ncu $w rc=$?"
done
