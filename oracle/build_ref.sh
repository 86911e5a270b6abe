#!/bin/bash
# oracle/build_ref.sh — TEST INFRASTRUCTURE.
# Compiles the UNMODIFIED reference gmove path from the sources where they lie under
# $REF (default /root/reference) into oracle/_ref/ (git-ignored, travels with gpurun).
# Nothing is copied from the reference into the repo; nothing is written under $REF.
#   oracle/_ref/libslow5.a      slow5lib + streamvbyte (reference submodule, C99, -O2 as its Makefile)
#   oracle/_ref/poregen_ref     src/gmove.cpp + poregen.cpp + thread.c, -O2
#   oracle/_ref/poregen_ref_g   same, flags as shipped by the reference Makefile:4 (-g, no -O)
#   oracle/_ref/reform_ref      src/reform.cpp (text SAM input only)
#   oracle/_ref/libsvbref.so    slow5lib's own svb-zd compress/depress behind two C entry points (svb_ref_wrap.c)
# htslib is not vendored by the reference (Makefile:69-81 downloads 1.17): the
# stand-in headers in oracle/shim/htslib are used instead (parsing only).
# No -march=native / -ffast-math: keeps fp64 results FMA-free (x86-64 baseline).
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${REF:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -d "$REF/src" ]; then
    echo "build_ref: $REF not present; keeping prebuilt files in $OUT (if any)"; exit 0
fi
mkdir -p "$OUT/obj"
S5="$REF/slow5lib"
SVB="$S5/thirdparty/streamvbyte"
CC=${CC:-gcc}; CXX=${CXX:-g++}
for f in streamvbyte_decode streamvbyte_encode streamvbyte_zigzag; do
    $CC -fPIC -std=c99 -O3 -DSTREAMVBYTE_SSSE3=1 -mssse3 -I "$SVB/include" -c "$SVB/src/$f.c" -o "$OUT/obj/$f.o"
done
for f in slow5 slow5_idx slow5_misc slow5_press; do
    $CC -g -O2 -std=c99 -fpic -I "$S5/include" -I "$SVB/include" -c "$S5/src/$f.c" -o "$OUT/obj/$f.o" 2>/dev/null
done
rm -f "$OUT/libslow5.a"
ar rcs "$OUT/libslow5.a" "$OUT"/obj/slow5*.o "$OUT"/obj/streamvbyte*.o
build_one() { # $1 = output name, rest = CFLAGS
    local name=$1; shift
    for f in gmove.cpp poregen.cpp thread.c; do
        $CXX -x c++ -std=c++11 "$@" -w -I "$S5/include" -I "$HERE/shim" -c "$REF/src/$f" -o "$OUT/obj/${name}_${f%.*}.o"
    done
    $CXX "$@" -c "$HERE/ref_main.cpp" -o "$OUT/obj/${name}_main.o"
    $CXX "$@" "$OUT/obj/${name}_main.o" "$OUT/obj/${name}_gmove.o" "$OUT/obj/${name}_poregen.o" "$OUT/obj/${name}_thread.o" \
        "$OUT/libslow5.a" -lz -lm -lpthread -o "$OUT/$name"
}
# `reform` (src/reform.cpp) against the same stand-in headers: text SAM only (the stand-in cannot open BAM).
cat > "$OUT/obj/reform_main.cpp" <<'EOM'
int reform(int argc, char *argv[]);
int main(int argc, char *argv[]) { return reform(argc - 1, argv + 1); }   /* reform_ref reform [options] file, like main.c */
EOM
for f in reform.cpp poregen.cpp thread.c; do
    $CXX -x c++ -std=c++11 -g -O2 -w -I "$S5/include" -I "$HERE/shim" -c "$REF/src/$f" -o "$OUT/obj/reform_${f%.*}.o"
done
$CXX -g -O2 "$OUT/obj/reform_main.cpp" "$OUT/obj/reform_reform.o" "$OUT/obj/reform_poregen.o" "$OUT/obj/reform_thread.o" "$OUT/libslow5.a" \
    -lz -lm -lpthread -o "$OUT/reform_ref"
$CC -g -O2 -std=c99 -fPIC -shared -I "$S5/include" "$HERE/svb_ref_wrap.c" "$OUT/libslow5.a" -lz -lm -o "$OUT/libsvbref.so"
build_one poregen_ref -g -O2
build_one poregen_ref_g -g
rm -rf "$OUT/obj"
echo "build_ref: built $OUT/poregen_ref $OUT/poregen_ref_g $OUT/reform_ref $OUT/libslow5.a"
