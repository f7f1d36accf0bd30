#!/bin/bash
# Build an experimental libmtscuda variant for kernel tuning sweeps:  tools/build_variant.sh NAME "TU TU ..." -DFOO=1 ...
# TU = shade translation units to recompile with the extra flags (c1p2 = k_shade_surface, c1p3 = k_shade_scatter, c1p0, c0p0,
# c2p0 = k_shade<class>, or `mtscuda` for the trace / raygen / splat kernels); everything else is linked from the regular build.
# The result (build/libmtscuda_NAME.so, git-ignored) is selected at run time with MTS_CUDA_LIB.
set -e
cd "$(dirname "$0")/../mitsuba0.6_b200/csrc"
name=$1; tus=$2; shift 2
mkdir -p ../../build
ARCH="-gencode arch=compute_100a,code=sm_100a"
FLAGS="$ARCH -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-fvisibility=hidden --expt-relaxed-constexpr -ccbin /usr/bin/g++"
objs=""
for o in mtscuda bvh_build shade_c0p0 shade_c0p1 shade_c1p0 shade_c1p1 shade_c1p2 shade_c1p3 shade_c1p4 shade_c2p0 shade_c2p1; do
  hit=0
  for t in $tus; do
    if [ "$o" = "$t" ] || [ "$o" = "shade_$t" ]; then hit=1; fi
  done
  if [ $hit = 1 ]; then
    out=../../build/${o}_$name.o
    if [ "$o" = "mtscuda" ]; then nvcc $FLAGS "$@" -c mtscuda.cu -o $out
    elif [ "$o" = "bvh_build" ]; then nvcc $FLAGS "$@" -c bvh_build.cu -o $out
    else
      cp=${o#shade_c}; cls=${cp%p*}; part=${cp#*p}
      nvcc $FLAGS "$@" -DSHADE_CLASS=$cls -DSHADE_PART=$part -c shade_inst.cu -o $out
    fi
    objs="$objs $out"
  else objs="$objs $o.o"; fi
done
nvcc $ARCH -shared -o ../../build/libmtscuda_$name.so $objs -lcudart_static -ldl -lrt -lpthread
rm -f ../../build/*_$name.o
echo built build/libmtscuda_$name.so
