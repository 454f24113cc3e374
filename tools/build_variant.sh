#!/bin/bash
# build_variant.sh <name> <file.cu> <extra nvcc flags...>: lib/variants/lib_<name>.so with one object rebuilt with flags
set -e
cd "$(dirname "$0")/../compactor-vllm_b200"
name=$1; src=$2; shift 2
mkdir -p lib/variants build/variants
obj=build/variants/${name}_$(basename $src .cu).o
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo --expt-relaxed-constexpr -Xcompiler -fPIC "$@" -c csrc/$src -o $obj
objs=$(ls build/*.o | grep -v "/$(basename $src .cu).o")
nvcc -shared -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -o lib/variants/lib_${name}.so $objs $obj
echo lib/variants/lib_${name}.so
