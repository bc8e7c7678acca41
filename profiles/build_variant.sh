#!/bin/bash
# build an experimental variant of the library: build_variant.sh NAME [-DFLAG ...]
# -> profiles/microbench/libb200head_NAME.so; select it with B200HEAD_LIB=<path> (see _lib.py)
set -e
name=$1; shift
cd "$(dirname "$0")/../mil-nce_howto100m_b200/csrc"
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -shared -Xcompiler -fPIC "$@" \
    -o ../../profiles/microbench/libb200head_$name.so capi.cu milnce.cu sdtw.cu
