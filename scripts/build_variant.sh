#!/bin/bash
# Kernel experiments: rebuild ONE source with extra -D flags and link it with the other objects of the regular build
# into build/lib<name>.so (select it with LG_B200_LIB=build/lib<name>.so).   usage: build_variant.sh name file.cu "-DX -DY"
set -e
cd "$(dirname "$0")/../cs566-project-lightglue_b200/csrc"
name=$1; src=$2; flags=$3
mkdir -p ../../build
/usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -I../../include -I. \
  --expt-relaxed-constexpr $flags -c $src -o ../../build/${name}_${src%.cu}.o
objs=$(ls lg_*.o | grep -v "^${src%.cu}.o$")
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../build/lib${name}.so $objs ../../build/${name}_${src%.cu}.o -lcudart
echo built build/lib${name}.so
