#!/bin/sh
# Build libhfrep.so in-tree for sm_100a (B200).  Usage: sh build.sh [extra nvcc flags]
# fit_kernel.cu is compiled with -fmad=false: its sums follow the canonical order, the only fused
# operations are the explicit fma intrinsics.  (Same recipe as __graft_entry__.build().)
set -e
here="$(cd "$(dirname "$0")" && pwd)"
out="$here/../libhfrep.so"
mkdir -p "$here/build"
flags="-std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC"
for f in common fit_kernel fit_cluster eval_kernel gather_kernel wide_kernel gen_kernel gen_tc; do
    extra=""
    case "$f" in fit_kernel|fit_cluster) extra="-fmad=false";; esac
    nvcc $flags $extra "$@" -c "$here/$f.cu" -o "$here/build/$f.o" &
done
wait
nvcc -shared -gencode arch=compute_100a,code=sm_100a "$here"/build/*.o -o "$out"
echo "built $out"
