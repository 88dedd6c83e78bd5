#!/bin/bash
# Build timing-experiment variants of libbfk in parallel: tools/build_variants.sh "name:flags" ...  -> build/libbfk_<name>.so
# (dev build: only the mish / D<=2 / fp16 instantiations).  Run on the GPU box with BFK_LIB=build/libbfk_<name>.so.
mkdir -p build
cd bayesflow_b200/csrc
for spec in "$@"; do
  name=${spec%%:*}; flags=${spec#*:}
  ( nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --shared -Xcompiler -fPIC -DBFK_DEV_BUILD $flags \
      -o ../../build/libbfk_$name.so bfk_host.cu > ../../build/$name.log 2>&1; echo "$name rc=$?" ) &
done
wait
