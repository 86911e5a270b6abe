#!/bin/bash
# Builds the `poregen gmove` command line of the B200 build: poregen_b200/bin/poregen.
# slow5lib (github.com/hasindu2008/slow5lib, the reference's vendored submodule) is an external
# dependency of the host side: it is compiled unchanged from $SLOW5LIB (default: the copy next to the
# reference) into poregen_b200/_deps/libslow5.a. Nothing of it is copied into this repository.
# Where $SLOW5LIB is absent (the GPU box) the prebuilt binary that travelled with the tree is kept.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
PKG="$(cd "$HERE/../.." && pwd)"
SLOW5LIB="${SLOW5LIB:-/root/reference/slow5lib}"
OUT="$PKG/bin/poregen"
if [ ! -d "$SLOW5LIB/src" ]; then
    echo "build_cli: $SLOW5LIB not present; keeping prebuilt $OUT (if any)"; exit 0
fi
if [ -x "$OUT" ] && [ "$OUT" -nt "$HERE/gmove_cli.cpp" ] && [ "$OUT" -nt "$HERE/bam_reader.hpp" ] && [ "$OUT" -nt "$HERE/reform.inc" ] && [ "$OUT" -nt "$PKG/../include/pgm_gmove.h" ] && [ "$OUT" -nt "$PKG/libpgm_gmove.so" ]; then
    exit 0
fi
CC=${CC:-gcc}; CXX=${CXX:-g++}
DEPS="$PKG/_deps"; mkdir -p "$DEPS/obj" "$PKG/bin"
SVB="$SLOW5LIB/thirdparty/streamvbyte"
if [ ! -f "$DEPS/libslow5.a" ]; then
    for f in streamvbyte_decode streamvbyte_encode streamvbyte_zigzag; do
        $CC -fPIC -std=c99 -O3 -DSTREAMVBYTE_SSSE3=1 -mssse3 -I "$SVB/include" -c "$SVB/src/$f.c" -o "$DEPS/obj/$f.o"
    done
    for f in slow5 slow5_idx slow5_misc slow5_press; do
        $CC -O2 -std=c99 -fpic -w -I "$SLOW5LIB/include" -I "$SVB/include" -c "$SLOW5LIB/src/$f.c" -o "$DEPS/obj/$f.o"
    done
    ar rcs "$DEPS/libslow5.a" "$DEPS"/obj/*.o
fi
$CXX -O2 -g -std=c++17 -Wall -I "$SLOW5LIB/include" "$HERE/gmove_cli.cpp" "$DEPS/libslow5.a" \
    -L "$PKG" -lpgm_gmove -Wl,-rpath,'$ORIGIN/..' -lz -lpthread -o "$OUT"
echo "build_cli: built $OUT"
