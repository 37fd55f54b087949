#!/bin/sh
# Builds libmbeast_sm100.so in-tree (the .so travels to the GPU box with the repo snapshot).
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/../libmbeast_sm100.so"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
"$NVCC" -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 \
    -Xcompiler -fPIC -Xcompiler -fvisibility=hidden -shared ${MB_NVCC_EXTRA} \
    -o "$OUT" "$HERE/abi.cu" "$HERE/head.cu" "$HERE/vtrace.cu" "$HERE/unpack.cu" "$HERE/convgrad.cu" "$HERE/gemm.cu" -lcudart_static -ldl -lrt -lpthread
echo "built $OUT"
