#!/bin/bash
# A/B builds of single kernels: tools/build_variant.sh NAME file.cu "-DFLAG=..."  ->  nerf_emitter_b200/libnerfemitter_NAME.so
# (all other objects are the ones of the normal build; select with NEMIT_LIB=nerf_emitter_b200/libnerfemitter_NAME.so)
set -e
cd "$(dirname "$0")/../nerf_emitter_b200/csrc"
NAME=$1; SRC=$2; FLAGS=$3
make -j8 > /dev/null
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xptxas -v --expt-relaxed-constexpr $FLAGS -c $SRC -o /tmp/variant_$NAME.o 2> /tmp/variant_$NAME.log
grep -E "registers|spill" /tmp/variant_$NAME.log | tail -4
OBJS=$(ls *.o | grep -v "^${SRC%.cu}.o$")
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../libnerfemitter_$NAME.so $OBJS /tmp/variant_$NAME.o -lcudart -ldl
echo built ../libnerfemitter_$NAME.so
