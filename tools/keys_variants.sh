#!/bin/bash
# Builds variants of csrc/keys2.cu (macros below) into csrc/var/libb2asr_<name>.so next to the objects of the last
# full build, for A/B timing on the GPU box:  B2ASR_LIB=$PWD/end-to-end-asr-pytorch_b200/csrc/var/libb2asr_<name>.so
# usage: tools/keys_variants.sh name="-DK2_STAGES=3" other="-DK2_NOCONV" ...
set -e
cd "$(dirname "$0")/../end-to-end-asr-pytorch_b200/csrc"
mkdir -p var
others=$(ls *.o | grep -v '^keys2\.o$')
for spec in "$@"; do
  name=${spec%%=*}; flags=${spec#*=}
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -DB2ASR_ARCH=100 \
       --expt-relaxed-constexpr $flags -c keys2.cu -o var/keys2_$name.o
  nvcc -shared -o var/libb2asr_$name.so $others var/keys2_$name.o -gencode arch=compute_100a,code=sm_100a -lcudart -lcublas
  echo "built var/libb2asr_$name.so ($flags)"
done
