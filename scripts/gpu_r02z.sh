#!/bin/bash
set -u
run() { echo "=== $*"; ( timeout 300 env "$@" ) 2>&1 | grep -v "^\[W" | tail -${TAILN:-3}; }
run X=1 python scripts/time_mixed.py
run B200ATTN_SORT_EXTEND=0 python scripts/time_mixed.py
run B200ATTN_SORT_EXTEND=0 B200ATTN_EXTEND_KV_SPLITS=1 python scripts/time_mixed.py
timeout 900 python -m pytest tests/test_gpu_extend.py tests/test_gpu_contract.py tests/test_gpu_masked.py -m gpu -x -q 2>&1 | tail -2
