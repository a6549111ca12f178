#!/bin/sh
# Development aid only: the cuemu build (see build.sh) with AddressSanitizer + UndefinedBehaviorSanitizer, for fuzzing the
# decoders' parsers on the CPU (compute-sanitizer is closed on the GPU pool).  Output: build/emu_san/ (never ships).
set -e
ROOT=$(cd "$(dirname "$0")/../.." && pwd)
OUT=$ROOT/build/emu_san
mkdir -p "$OUT"
SAN="-fsanitize=address,undefined -fno-sanitize-recover=undefined -fno-omit-frame-pointer"
g++ -O1 -g $SAN -fPIC -ffp-contract=off -std=c++17 -w -I"$ROOT/tools/cuemu" -include "$ROOT/tools/cuemu/cuemu.h" \
    -x c++ -c "$ROOT/ebcc_b200/csrc/ebcc_cuda.cu" -o "$OUT/ebcc_cuda.o"
g++ -O1 -g $SAN -fPIC -std=c++17 -w -c "$ROOT/tools/cuemu/cuemu.cpp" -o "$OUT/cuemu.o"
for f in ebcc_codec h5z_ebcc; do
    gcc -O1 -g $SAN -fPIC -ffp-contract=off -w -c "$ROOT/ebcc_b200/csrc/$f.c" -o "$OUT/$f.o"
done
g++ -shared $SAN -o "$OUT/libh5z_ebcc_emu_san.so" "$OUT/ebcc_cuda.o" "$OUT/cuemu.o" "$OUT/ebcc_codec.o" "$OUT/h5z_ebcc.o" -lpthread -lm
echo "$OUT/libh5z_ebcc_emu_san.so"
