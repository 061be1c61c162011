#!/bin/bash
# build_variant.sh <tag> <source.cu> <nvcc -D flags...>: libmdv2_<tag>.so = the current objects with ONE source
# recompiled under extra flags (same-box A/B through MDV2_LIB_FILE)
set -e
cd "$(dirname "$0")/../magicdrive_v2_b200"
tag=$1; src=$2; shift 2
obj=build/${src%.cu}_$tag.o
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC \
  --expt-relaxed-constexpr "$@" -c csrc/$src -o $obj
objs=""
for o in build/*.o; do
  b=$(basename $o .o)
  case "$b" in
    ${src%.cu}) ;;                     # replaced
    *_v[0-9a-z]*) ;;                   # other variants
    *) objs="$objs $o";;
  esac
done
/usr/local/cuda/bin/nvcc -shared -o libmdv2_$tag.so $objs $obj -gencode arch=compute_100a,code=sm_100a -lcudart
echo built libmdv2_$tag.so
