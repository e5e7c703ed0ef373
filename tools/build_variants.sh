#!/bin/bash
# usage: bash tools/build_variants.sh name1 "flags1" name2 "flags2" ...
# Builds monodepth2.jl_b200/variants/libmdb200_<name>.so: photo_fused.cu recompiled with the extra -D flags, every other
# object taken from the standard build.  Done HERE (nvcc cross-compiles); tools/time_variants.py times them on the GPU box.
set -e
cd "$(dirname "$0")/../monodepth2.jl_b200"
make -j8 >/dev/null
mkdir -p variants
ARCH="-gencode arch=compute_100a,code=sm_100a"
while [ $# -gt 1 ]; do
  name=$1; flags=$2; shift 2
  (
    mkdir -p build_var/$name
    nvcc $ARCH -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden --expt-relaxed-constexpr $flags -Xptxas -v \
      -c csrc/photo_fused.cu -o build_var/$name/photo_fused.o 2> build_var/$name/ptxas.log
    others=$(ls build/*.o | grep -v photo_fused.o)
    nvcc $ARCH -shared -o variants/libmdb200_$name.so build_var/$name/photo_fused.o $others -lcudart
    echo "$name: $flags :: $(grep -A2 'photo_strip_kernelILi3ELi2ELb0' build_var/$name/ptxas.log | grep -E 'spill|Used' | tr '\n' ' ')"
  ) &
done
wait
