#!/bin/bash
# build_ab/libptb200_<name>.so: the library with extra -D flags on the two kernel translation units (A/B experiments, PTB200_LIB selects it)
#   scripts/build_variant.sh ns40 -DPOOL_NS=40
set -e
name=$1; shift
cd "$(dirname "$0")/../opencl_path_tracer_b200/csrc"
mkdir -p ../../build_ab/obj_$name build
ARCH="-gencode arch=compute_100a,code=sm_100a"
COMMON="-O3 -std=c++17 -lineinfo -Xcompiler -fPIC -I. -I../../include $ARCH"
nvcc $COMMON --fmad=true -Xptxas -v "$@" -c pt_kernels_fast.cu -o ../../build_ab/obj_$name/pt_kernels_fast.o 2> ../../build_ab/obj_$name/ptxas_fast.log &
nvcc $COMMON --fmad=false --prec-div=true --prec-sqrt=true --ftz=false "$@" -c pt_kernels_parity.cu -o ../../build_ab/obj_$name/pt_kernels_parity.o 2> ../../build_ab/obj_$name/ptxas_parity.log &
wait
nvcc $ARCH -shared -o ../../build_ab/libptb200_$name.so build/ptb_abi.o ../../build_ab/obj_$name/pt_kernels_parity.o ../../build_ab/obj_$name/pt_kernels_fast.o \
    build/lbvh.o build/resolve.o build/prims.o build/ingest.o build/comm.o -lcudart_static -lpthread -ldl -lrt
grep -A2 "k_poolILb0ELi0" ../../build_ab/obj_$name/ptxas_fast.log | grep -E "Used|spill" | head -3
