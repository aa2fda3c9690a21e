#!/bin/bash
# build_variant.sh <name> <file.cu> <extra nvcc flags...>: liborbx with ONE translation unit rebuilt with extra flags
# -> .variants/liborbx_<name>.so (git-ignored; used with ORBX_LIB=... for kernel A/B runs on the GPU box)
set -e
cd "$(dirname "$0")/.."
name=$1; src=$2; shift 2
mkdir -p .variants/obj
base=$(basename $src .cu)
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-ffp-contract=off,-Wall "$@" -c -o .variants/obj/${base}_$name.o orb-slam2-noted_b200/csrc/$src
objs=$(ls orb-slam2-noted_b200/build/*.o | grep -v "/$base.o")
/usr/local/cuda/bin/nvcc -shared -o .variants/liborbx_$name.so $objs .variants/obj/${base}_$name.o -gencode arch=compute_100a,code=sm_100a
echo .variants/liborbx_$name.so
