#!/bin/sh
# Builds libmbeast_sm100.so in-tree (the .so travels to the GPU box with the repo snapshot).
# Each translation unit is compiled to build/<name>.o, in parallel and only when it is older than its
# sources, then linked; MB_NVCC_EXTRA adds flags (e.g. "-Xptxas -v"), MB_REBUILD=1 forces everything.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/../libmbeast_sm100.so"
OBJ="$HERE/build"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden ${MB_NVCC_EXTRA}"
mkdir -p "$OBJ"
UNITS="abi head vtrace unpack convgrad convtc gemm"
[ -f "$HERE/cells.cu" ] && UNITS="$UNITS cells"
[ -f "$HERE/fused.cu" ] && UNITS="$UNITS fused"
[ -f "$HERE/sampler.cu" ] && UNITS="$UNITS sampler"
[ -f "$HERE/convbits.cu" ] && UNITS="$UNITS convbits"
pids=""
for u in $UNITS; do
    src="$HERE/$u.cu"
    obj="$OBJ/$u.o"
    if [ -n "$MB_REBUILD" ] || [ -n "$MB_NVCC_EXTRA" ] || [ ! -f "$obj" ] || [ "$src" -nt "$obj" ] || \
       [ "$HERE/common.cuh" -nt "$obj" ] || [ "$HERE/tc.cuh" -nt "$obj" ] || [ "$HERE/../../include/mbeast.h" -nt "$obj" ] || [ "$0" -nt "$obj" ]; then
        rm -f "$obj"
        ( "$NVCC" $FLAGS -c -o "$obj" "$src" 2>&1 | sed "s/^/[$u] /" ; ) &
        pids="$pids $!"
    fi
done
for p in $pids; do wait "$p"; done
objs=""
for u in $UNITS; do
    [ -f "$OBJ/$u.o" ] || { echo "build.sh: $u.cu failed to compile" >&2; exit 1; }
    objs="$objs $OBJ/$u.o"
done
"$NVCC" -gencode arch=compute_100a,code=sm_100a -shared -o "$OUT" $objs -lcudart_static -ldl -lrt -lpthread
echo "built $OUT"
