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
    # no sources to rebuild from (the GPU box): the prebuilt binary must match the sources that travelled with it
    SRC_HASH="$(cat "$HERE/gmove_cli.cpp" "$HERE/bam_reader.hpp" "$HERE/reform.inc" "$PKG/../include/pgm_gmove.h" | sha256sum | cut -c1-64)"
    if [ -x "$OUT" ] && [ "$("$OUT" --build-info 2>/dev/null || true)" != "$SRC_HASH" ]; then
        echo "build_cli: $OUT is STALE (built from other sources than the tree holds) and $SLOW5LIB is absent: cannot rebuild" >&2; exit 3
    fi
    echo "build_cli: $SLOW5LIB not present; keeping prebuilt $OUT (if any)"; exit 0
fi
SRC_HASH="$(cat "$HERE/gmove_cli.cpp" "$HERE/bam_reader.hpp" "$HERE/reform.inc" "$PKG/../include/pgm_gmove.h" | sha256sum | cut -c1-64)"
CLILIB="$PKG/libporegen_gmove_cli.so"
if [ -x "$OUT" ] && [ -f "$CLILIB" ] && [ "$("$OUT" --build-info 2>/dev/null || true)" = "$SRC_HASH" ] && [ "$OUT" -nt "$PKG/libpgm_gmove.so" ]; then
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
$CXX -O2 -g -std=c++17 -Wall -DPGM_SRC_HASH="\"$SRC_HASH\"" -I "$SLOW5LIB/include" "$HERE/gmove_cli.cpp" "$DEPS/libslow5.a" \
    -L "$PKG" -lpgm_gmove -Wl,-rpath,'$ORIGIN/..' -lz -lpthread -o "$OUT"
# the same code as a library exporting gmove(int, char**) / reform(int, char**) (include/pgm_gmove_cli.h): what the
# reference's own main.o links against
$CXX -O2 -g -std=c++17 -Wall -fPIC -shared -DPGM_NO_MAIN -DPGM_SRC_HASH="\"$SRC_HASH\"" -I "$SLOW5LIB/include" "$HERE/gmove_cli.cpp" "$DEPS/libslow5.a" \
    -L "$PKG" -lpgm_gmove -Wl,-rpath,'$ORIGIN' -lz -lpthread -o "$CLILIB"
echo "build_cli: built $OUT and $CLILIB"
