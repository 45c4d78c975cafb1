#!/bin/bash
# round 2, call H: the even-stage-count fix of the two-tile-group decode kernel -- failing shapes, parity, C4 timing
set -u
O=gpurun_out
mkdir -p $O
run() { echo "=== $*"; ( timeout 120 env "$@" ) 2>&1 | grep -v "^\[W" | tail -3; }
run X=1 python scripts/debug_paged.py 64 64 8 32000 1 16 148
run X=1 python scripts/debug_paged.py 16 64 8 32000 1 16 148
run X=1 python scripts/debug_paged.py 64 64 8 8000 1 16 148
run X=1 python scripts/debug_paged.py 248 64 8 0 1 16 148
run X=1 python scripts/debug_paged.py 248 32 4 0 1 16 148
run X=1 python scripts/debug_paged.py 248 16 2 0 1 16 148
timeout 600 python -m pytest tests/test_gpu_decode.py tests/test_gpu_tp.py -m gpu -x -q > $O/pytest_decode_r02h.log 2>&1; echo "pytest decode rc=$?"
tail -4 $O/pytest_decode_r02h.log
timeout 300 python scripts/sweep_shapes.py c4 2>&1 | tee $O/sweep_c4_r02h.txt
timeout 300 python scripts/sweep_shapes.py c4layout 2>&1 | tee $O/sweep_c4layout_r02h.txt
