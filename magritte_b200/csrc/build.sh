#!/bin/bash
# Build libmagritte_b200.so (hand-written CUDA for sm_100a + the C ABI of include/magritte_b200.h) in-tree.
# -fmad=false: the reference build contains no fused multiply-adds; every fma in the
# kernels is an explicit fma()/__fma_rn call. -lineinfo: ncu source pages map to these files.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/../lib"
mkdir -p "$OUT"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -Xcompiler -fPIC,-O2,-fno-fast-math -Xptxas -v"
[ -n "$MGB_NO_TEST_HOOKS" ] && FLAGS="$FLAGS -DMGB_NO_TEST_HOOKS"
pids=()
for f in mgb_api mgb_debug mgb_comm mgb_image mgb_levelpop mgb_noalloc; do
    if [ ! -f "$OUT/$f.o" ] || [ -n "$(find "$HERE" "$HERE/../../include" -newer "$OUT/$f.o" \( -name '*.cu' -o -name '*.cuh' -o -name '*.h' \) | head -1)" ]; then
        $NVCC $FLAGS -c "$HERE/$f.cu" -o "$OUT/$f.o" > "$OUT/$f.ptxas.log" 2>&1 &
        pids+=($!)
    fi
done
rc=0
for p in "${pids[@]}"; do wait "$p" || rc=1; done
if [ $rc -ne 0 ]; then cat "$OUT"/*.ptxas.log; exit 1; fi
$NVCC -shared -o "$OUT/libmagritte_b200.so" "$OUT/mgb_api.o" "$OUT/mgb_debug.o" "$OUT/mgb_comm.o" "$OUT/mgb_image.o" "$OUT/mgb_levelpop.o" "$OUT/mgb_noalloc.o" -lcudart -ldl
echo "built $OUT/libmagritte_b200.so"
