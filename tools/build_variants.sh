#!/bin/bash
# Build libduckhts_cuda.so variants of bgzf_inflate3.cu for A/B runs on the GPU box:
#   tools/build_variants.sh name1:"-DFOO=1 -DBAR=2" name2:"..."   -> variants/<name>/libduckhts_cuda.so  (git-ignored, travels with gpurun)
# The default library under duckhts_b200/lib/ must be built first (its other objects are reused).
set -e
cd "$(dirname "$0")/../duckhts_b200/csrc"
LIB=../lib
for spec in "$@"; do
  name="${spec%%:*}"; flags="${spec#*:}"
  out=../../variants/$name; mkdir -p $out
  nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC $flags -c -o $out/bgzf_inflate3.o bgzf_inflate3.cu
  objs=$(ls $LIB/obj/*.o | grep -v bgzf_inflate3.o)
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $out/libduckhts_cuda.so $objs $out/bgzf_inflate3.o -Xlinker -soname,libduckhts_cuda.so
  echo "built $out ($flags)"
done
