#!/bin/bash
# tools/build_variant.sh NAME 'sed-expr' ['extra nvcc flags'] [file-to-sed] : variant library lib/variants/libmgb_NAME.so with a sed edit applied to
# one source file (default mgb_solve.cuh) and/or extra -D flags (MGB_LIB=... selects it at run time)
set -e
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
SRC="$ROOT/magritte_b200/csrc"; OUT="$ROOT/magritte_b200/lib/variants"; T=$(mktemp -d)
mkdir -p "$OUT"; cp "$SRC"/*.cu "$SRC"/*.cuh "$SRC"/*.h "$T"/
sed -i "s#../../include/magritte_b200.h#$ROOT/include/magritte_b200.h#" "$T/mgb_ctx.h"
[ -n "$2" ] && sed -i -E "$2" "$T/${4:-mgb_solve.cuh}"
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -Xcompiler -fPIC,-O2,-fno-fast-math -Xptxas -v $3"
OBJS=""
for f in mgb_api mgb_debug mgb_comm mgb_image mgb_levelpop mgb_noalloc; do nvcc $FLAGS -c "$T/$f.cu" -o "$T/$f.o" > "$T/$f.log" 2>&1 & OBJS="$OBJS $T/$f.o"; done; wait
cp "$T/mgb_api.log" "$OUT/$1.ptxas.log"; grep -A2 "k_sweepILi2\|k_largeILi2\|k_fillEN" "$T/mgb_api.log" | grep -E "Used|spill" | tr '\n' ' '; echo
nvcc -shared -o "$OUT/libmgb_$1.so" $OBJS -lcudart -ldl 2>/dev/null
rm -rf "$T"
