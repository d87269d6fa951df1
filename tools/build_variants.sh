#!/bin/bash
# Kernel A/B builds: libse_b200.so with a different set of -D flags for the NoiseSup translation unit, under
# speech_enhancement_b200/lib/variants/<tag>/ (git-ignored; they travel to the GPU box with gpurun).  Select one at run
# time with SE_B200_LIB=<path>.   usage: tools/build_variants.sh tag "-DFLAG ..." [tag "-D..."]...
set -e
cd "$(dirname "$0")/../speech_enhancement_b200/csrc"
make -s -j4
while [ $# -ge 2 ]; do
  tag=$1; flags=$2; shift 2
  out=../lib/variants/$tag; mkdir -p $out build/variants
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC $flags -c se_ns.cu -o build/variants/se_ns_$tag.o
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $out/libse_b200.so build/variants/se_ns_$tag.o build/se_ns_exact.o build/se_g64.o build/se_api.o build/se_tables.o
  echo "built $out/libse_b200.so ($flags)"
done
