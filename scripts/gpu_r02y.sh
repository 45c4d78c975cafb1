#!/bin/bash
set -u
run() { echo "=== $*"; ( timeout 300 env "$@" ) 2>&1 | grep -v "^\[W" | tail -${TAILN:-4}; }
run X=1 python scripts/debug_extend.py 1 30000 2048 8 1 splits=6
run X=1 python scripts/debug_extend.py 1 30000 2048 32 8 splits=3
run X=1 python scripts/debug_extend.py 2 8192 1024 8 1 splits=4
TAILN=3 run X=1 python scripts/time_mixed.py
TAILN=3 run B200ATTN_EXTEND_KV_SPLITS=1 python scripts/time_mixed.py
timeout 900 python -m pytest tests/test_gpu_extend.py tests/test_gpu_contract.py tests/test_gpu_tp.py -m gpu -x -q 2>&1 | tail -2
