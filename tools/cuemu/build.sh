#!/bin/sh
# Development aid only: compiles the CUDA translation unit as host C++ against tools/cuemu/cuemu.h so
# that kernel control flow can be debugged in a container without a GPU.  Output goes to build/emu/
# (git-ignored AND gpurun-ignored): it never ships and the product never loads it.
set -e
ROOT=$(cd "$(dirname "$0")/../.." && pwd)
OUT=$ROOT/build/emu
mkdir -p "$OUT"
g++ -O2 -g -fPIC -ffp-contract=off -std=c++17 -w -I"$ROOT/tools/cuemu" -include "$ROOT/tools/cuemu/cuemu.h" \
    -x c++ -c "$ROOT/ebcc_b200/csrc/ebcc_cuda.cu" -o "$OUT/ebcc_cuda.o"
g++ -O2 -g -fPIC -std=c++17 -w -c "$ROOT/tools/cuemu/cuemu.cpp" -o "$OUT/cuemu.o"
for f in ebcc_codec h5z_ebcc; do
    gcc -O2 -g -fPIC -ffp-contract=off -w -c "$ROOT/ebcc_b200/csrc/$f.c" -o "$OUT/$f.o"
done
g++ -shared -o "$OUT/libh5z_ebcc_emu.so" "$OUT/ebcc_cuda.o" "$OUT/cuemu.o" "$OUT/ebcc_codec.o" "$OUT/h5z_ebcc.o" -lpthread -lm
echo "$OUT/libh5z_ebcc_emu.so"
