#!/bin/bash
# Build an experimental libmtscuda variant for kernel tuning sweeps:  tools/build_variant.sh NAME -DFOO=1 ...
# The result (build/libmtscuda_NAME.so, git-ignored) is selected at run time with MTS_CUDA_LIB.
set -e
cd "$(dirname "$0")/../mitsuba0.6_b200/csrc"
name=$1; shift
ARCH="-gencode arch=compute_100a,code=sm_100a"
nvcc $ARCH -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-fvisibility=hidden --expt-relaxed-constexpr -ccbin /usr/bin/g++ "$@" -c mtscuda.cu -o ../../build/mtscuda_$name.o
nvcc $ARCH -shared -o ../../build/libmtscuda_$name.so ../../build/mtscuda_$name.o bvh_build.o -lcudart_static -ldl -lrt -lpthread
rm -f ../../build/mtscuda_$name.o
echo built build/libmtscuda_$name.so
