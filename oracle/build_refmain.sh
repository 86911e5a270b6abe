#!/bin/bash
# oracle/build_refmain.sh - TEST INFRASTRUCTURE.
# Links the reference's UNMODIFIED dispatcher (src/main.c, compiled as C++ like its Makefile:5,39 does) against THIS
# build's libporegen_gmove_cli.so: the proof that `int gmove(int, char**)` / `int reform(int, char**)` are drop-in at
# the link level. Output: oracle/_ref/poregen_refmain (git-ignored, travels with gpurun).
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${REF:-/root/reference}"
PKG="$(cd "$HERE/../poregen_b200" && pwd)"
OUT="$HERE/_ref/poregen_refmain"
if [ ! -f "$REF/src/main.c" ]; then echo "build_refmain: $REF not present; keeping prebuilt $OUT (if any)"; exit 0; fi
if [ ! -f "$PKG/libporegen_gmove_cli.so" ]; then echo "build_refmain: $PKG/libporegen_gmove_cli.so missing" >&2; exit 1; fi
mkdir -p "$HERE/_ref"
CXX=${CXX:-g++}
$CXX -x c++ -std=c++11 -g -O2 -w -I "$REF/src" -I "$REF/slow5lib/include" -c "$REF/src/main.c" -o "$HERE/_ref/refmain_main.o"
$CXX -std=c++11 -g -O2 -w -I "$REF/src" -I "$REF/slow5lib/include" -c "$HERE/refmain_stubs.cpp" -o "$HERE/_ref/refmain_stubs.o"
$CXX "$HERE/_ref/refmain_main.o" "$HERE/_ref/refmain_stubs.o" -L "$PKG" -lporegen_gmove_cli -lpgm_gmove \
    -Wl,-rpath,'$ORIGIN/../../poregen_b200' -o "$OUT"
rm -f "$HERE/_ref/refmain_main.o" "$HERE/_ref/refmain_stubs.o"
echo "build_refmain: built $OUT"
