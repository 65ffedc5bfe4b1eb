#!/bin/sh
# Build libhfrep.so in-tree for sm_100a (B200).  Usage: sh build.sh [extra nvcc flags]
set -e
here="$(cd "$(dirname "$0")" && pwd)"
out="$here/../libhfrep.so"
nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo \
     -Xcompiler -fPIC -shared "$@" \
     "$here/common.cu" "$here/fit_kernel.cu" "$here/eval_kernel.cu" "$here/gather_kernel.cu" "$here/wide_kernel.cu" \
     -o "$out"
echo "built $out"
