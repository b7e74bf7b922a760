#!/bin/bash
# Development aid: build a variant of the CUDA library with extra -D flags into ktx_software_b200/variants/ (git-ignored *.so;
# select it with B200_UASTC_LIB=...).  usage: tools/build_variant.sh NAME [-DFOO=1 ...]
set -e
cd "$(dirname "$0")/.."
N=$1; shift
mkdir -p ktx_software_b200/variants
exec /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 --fmad=false -std=c++17 -Xcompiler -fPIC -shared "$@" \
  -o ktx_software_b200/variants/libuastc_$N.so ktx_software_b200/csrc/uastc_b200.cu
