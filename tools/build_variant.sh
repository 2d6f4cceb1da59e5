#!/bin/bash
# Builds a variant of the library with one source recompiled under extra flags (builder experiments; git-ignored output):
#   tools/build_variant.sh <name> <source.cu> <extra nvcc flags...>   ->  cosyvoice_b200/lib/libcvb200_<name>.so   (use with CVB200_LIB=...)
set -e
cd "$(dirname "$0")/../cosyvoice_b200"
name=$1; src=$2; shift 2
python -m cosyvoice_b200.build >/dev/null 2>&1 || (cd .. && python -m cosyvoice_b200.build >/dev/null)
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr "$@" -c csrc/$src -o build/variant_${name}.o
objs=$(ls build/*.o | grep -v variant_ | grep -v "/${src%.cu}.o")
/usr/local/cuda/bin/nvcc -shared -o lib/libcvb200_${name}.so $objs build/variant_${name}.o -lcudart
echo lib/libcvb200_${name}.so
