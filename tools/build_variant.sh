#!/bin/bash
# development aid: the product library with extra -D flags, as build/var/<name>.so (select it with EBCC_PROD_SO=...)
# usage: tools/build_variant.sh <name> [-DFLAG=VALUE ...]
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p build/var build/obj
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -fmad=false -std=c++17 -Xcompiler -fPIC "$@" -c ebcc_b200/csrc/ebcc_cuda.cu -o build/var/$name.o
nvcc -arch=sm_100a -shared -cudart static -o build/var/$name.so build/var/$name.o build/obj/ebcc_codec.o build/obj/h5z_ebcc.o -lpthread -lm
rm -f build/var/$name.o
ls -la build/var/$name.so
