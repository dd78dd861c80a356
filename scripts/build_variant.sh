#!/bin/bash
# Build an experimental copy of librrt.so with extra nvcc flags:
#   scripts/build_variant.sh NAME [-DRRT_MIN_BLOCKS=5 ...]   -> rustraytrace_b200/variants/librrt_NAME.so
# Select it at run time with RRT_LIB=variants/librrt_NAME.so (rustraytrace_b200/_ffi.py).
set -e
cd "$(dirname "$0")/.."
name=$1; shift
out=rustraytrace_b200/variants; tmp=$(mktemp -d)
mkdir -p $out
C=rustraytrace_b200/csrc
F="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC"
nvcc $F --use_fast_math "$@" -c $C/kernels.cu -o $tmp/kernels.o &
for b in 1 2 3; do nvcc $F --use_fast_math "$@" -c $C/kernels_book$b.cu -o $tmp/kernels_book$b.o & done
nvcc $F -fmad=false "$@" -c $C/probe64.cu -o $tmp/probe64.o &
nvcc $F --use_fast_math "$@" -c $C/api.cu -o $tmp/api.o &
nvcc $F --use_fast_math "$@" -c $C/multi.cu -o $tmp/multi.o &
nvcc $F --use_fast_math "$@" -c $C/noalloc.cu -o $tmp/noalloc.o &
nvcc $F --use_fast_math "$@" -x cu -c $C/flatten.cpp -o $tmp/flatten.o &
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $out/librrt_$name.so $tmp/kernels.o $tmp/kernels_book1.o $tmp/kernels_book2.o $tmp/kernels_book3.o $tmp/probe64.o $tmp/api.o $tmp/multi.o $tmp/noalloc.o $tmp/flatten.o -cudart shared -lnccl
rm -rf $tmp
echo built $out/librrt_$name.so
