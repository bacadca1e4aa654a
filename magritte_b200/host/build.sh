#!/bin/bash
# Build the host layer: magritte_b200/core*.so = pybind11 module over mgbh::Model (C++17, g++), linked against the C-ABI library.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
PKG="$HERE/.."
PY="${PYTHON:-python}"
EXT=$($PY -c "import sysconfig; print(sysconfig.get_config_var('EXT_SUFFIX'))")
INC=$($PY -c "import pybind11, sysconfig; print('-I' + pybind11.get_include() + ' -I' + sysconfig.get_paths()['include'])")
OUT="$PKG/core$EXT"
if [ ! -f "$OUT" ] || [ "$HERE/bindings.cpp" -nt "$OUT" ] || [ "$HERE/mgb_host.hpp" -nt "$OUT" ] || [ "$PKG/../include/magritte_b200.h" -nt "$OUT" ]; then
    g++ -O3 -DNDEBUG -std=c++17 -fopenmp -fPIC -shared -fvisibility=hidden -fno-fast-math -ffp-contract=off $INC "$HERE/bindings.cpp" -o "$OUT" \
        -L"$PKG/lib" -lmagritte_b200 -Wl,-rpath,'$ORIGIN/lib'
fi
echo "built $OUT"
