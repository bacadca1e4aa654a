#!/bin/bash
# Compile the UNMODIFIED reference (Magritte v0.9.3) from the sources where they lie
# under /root/reference into oracle/_ref/ (git-ignored). TEST INFRASTRUCTURE ONLY:
# the result is the parity oracle and the timed CPU baseline, never the product.
#
# The reference's own CMake build cannot configure here (CMakeLists.txt:38 forces
# -std=c++17 into CMAKE_C_FLAGS, FindOpenMP for C then fails), so this is the direct
# g++ loop over the 23 sources of src/bindings/CMakeLists.txt:8-32 with the Release
# flags of CMakeLists.txt:28,37 (-O3 -DNDEBUG -std=c++17 -fopenmp; no -march, no
# -ffast-math). The two generated configure.hpp headers are replaced by the
# hand-written ones under oracle/ref_cfg (resolved through -I oracle/ref_cfg/a/b).
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${MAGRITTE_REFERENCE:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -d "$REF/src" ]; then
    echo "build_ref: $REF not present; keeping prebuilt $OUT (if any)"; exit 0
fi
mkdir -p "$OUT/obj" "$OUT/magritte"
PYINC="$(python3 -c 'import sysconfig; print(sysconfig.get_paths()["include"])')"
[ -f "$PYINC/Python.h" ] || PYINC=/usr/include/python3.12
PYLIBDIR="$(python3 -c 'import sysconfig; print(sysconfig.get_config_var("LIBDIR"))')"
CXXFLAGS="-O3 -DNDEBUG -std=c++17 -fopenmp -fPIC -fvisibility=hidden -w"
INCS="-I$REF/src -I$REF/dependencies -I$REF/dependencies/eigen -I$REF/dependencies/pybind11/include \
 -I$REF/dependencies/Paracabs/Paracabs -I$HERE/ref_cfg/a/b -I$PYINC"
SRCS=$(sed -n '8,32p' "$REF/src/bindings/CMakeLists.txt" | tr -d ' ' | grep cpp)
pids=()
for s in $SRCS; do
    src="$REF/src/bindings/$s"
    obj="$OUT/obj/$(basename "$s" .cpp).o"
    if [ ! -f "$obj" ] || [ "$src" -nt "$obj" ]; then
        g++ $CXXFLAGS '-DVERSION_NUMBER="0.9.3"' $INCS -c "$src" -o "$obj" &
        pids+=($!)
    fi
done
for p in "${pids[@]}"; do wait "$p"; done
g++ -shared -fopenmp "$OUT"/obj/*.o -o "$OUT/magritte/core.so" -L"$PYLIBDIR" -lpython3.12
touch "$OUT/magritte/__init__.py"
# Trace hook (our own file, compiled against the reference headers; exports the private
# per-thread scratch nr_/shift_/dZ_ of Solver::trace_ray_for_line for the bit-exact check).
if [ -f "$HERE/ref_hooks.cpp" ]; then
    g++ $CXXFLAGS '-DVERSION_NUMBER="0.9.3"' $INCS -c "$HERE/ref_hooks.cpp" -o "$OUT/obj_ref_hooks.o"
    objs=$(ls "$OUT"/obj/*.o | grep -v pybindings.o)
    g++ -shared -fopenmp $objs "$OUT/obj_ref_hooks.o" -o "$OUT/magritte/ref_hooks.so" -L"$PYLIBDIR" -lpython3.12
fi
echo "build_ref: ok -> $OUT/magritte/core.so"
