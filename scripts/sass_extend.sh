#!/bin/bash
# Compile csrc/extend_tc.cu alone (sm_100a, -Xptxas -v) and dump the SASS of one kernel instance to /tmp/k3.txt.
# usage: scripts/sass_extend.sh [mangled-name-substring]   (default: bf16, kEmu = 2, no prof)
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
SRC=$ROOT/sglang_custom_b200/csrc/extend_tc.cu
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --expt-relaxed-constexpr -Xcompiler -fPIC \
  -Xcompiler -fvisibility=hidden -I "$ROOT/include" -Xptxas -v -c "$SRC" -o /tmp/extend_tc.o 2>&1 | grep -E "error|spill|Used" | head -${LINES_V:-4}
FUN=${1:-_ZN8b200attn2tc16extend_tc_kernelILb1ELi2ELb0EEEvNS_12ExtendParamsEi}
cuobjdump -sass -fun "$FUN" /tmp/extend_tc.o | grep -E "^\s+/\*[0-9a-f]{4}\*/" | sed -E 's/^\s+\/\*([0-9a-f]+)\*\/\s+/\1 /; s/\s*\/\*.*//' > /tmp/k3.txt
echo "LDL/STL at lines:"; grep -n "USETMAXREG\|LDL\|STL" /tmp/k3.txt | awk -F: '{printf "%s ", $1}'; echo; wc -l /tmp/k3.txt
