#!/bin/bash
set -u
run() { echo "=== $*"; ( timeout 300 env "$@" ) 2>&1 | grep -v "^\[W" | tail -${TAILN:-5}; }
run X=1 python scripts/debug_extend.py 1 30000 2048 8 1 splits=4
run X=1 python scripts/debug_extend.py 2 3000 700 ragged splits=2
run X=1 python scripts/debug_extend.py 3 100 300 ragged splits=4
run X=1 python scripts/debug_extend.py 2 8000 1024 8 1 fp8 splits=3
run X=1 python scripts/debug_extend.py 2 8192 1024 8 1 page16 splits=2
run X=1 python scripts/debug_extend.py 1 20000 1500 28 4 splits=2
run X=1 python scripts/debug_extend.py 8 12288 1024 8 1 splits=2
