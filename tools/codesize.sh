#!/bin/bash
# Development aid: compile the CUDA module to a cubin and list per-function SASS sizes (bytes) and register counts.
set -e
cd "$(dirname "$0")/.."
OUT=${TMPDIR:-/tmp}/uastc_b200.cubin
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 --fmad=false -std=c++17 -cubin -o $OUT ktx_software_b200/csrc/uastc_b200.cu "$@"
readelf -sW $OUT 2>/dev/null | awk '$4=="FUNC"{print $3, $8}' | sed 's/\$_ZN46_GLOBAL__N__[a-z0-9_]*_cu_[a-z0-9]*//; s/_ZN46_GLOBAL__N__[a-z0-9_]*_cu_[a-z0-9]*//' | sort -n | tail -${TAILN:-40}
cuobjdump -res-usage $OUT | grep -A1 "Function" | grep -o "Function [^:]*\|REG:[0-9]* STACK:[0-9]*" | paste - - | sed 's/_ZN46_GLOBAL__N__[a-z0-9_]*_cu_[a-z0-9]*//'
