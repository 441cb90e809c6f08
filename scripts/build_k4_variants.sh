#!/bin/bash
# Variant builds of the library for an A/B on one box (scripts/r02_ab.sh): only nb_likelihood_fast.cu is recompiled with the
# variant's -D flags, the other objects are the default build's.  usage: build_k4_variants.sh name:"flags" ...
set -e
cd "$(dirname "$0")/../mm-vae_b200"
make -j8 > /dev/null
mkdir -p ../gpurun_libs
cp libmmvae_b200.so ../gpurun_libs/base.so
for spec in "$@"; do
  name=${spec%%:*}; flags=${spec#*:}
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden $flags \
       -c csrc/nb_likelihood_fast.cu -o /tmp/k4_$name.o
  objs=$(ls build/*.o | grep -v nb_likelihood_fast.o)
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../gpurun_libs/$name.so $objs /tmp/k4_$name.o -ldl
  echo "built gpurun_libs/$name.so ($flags)"
done
