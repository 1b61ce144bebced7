#!/bin/bash
# Build an A/B variant of libb2conv.so: tools/build_variant.sh NAME "-DFLAG=1 ..." [source.cu]  -> build/variants/NAME.so
# (the library must have been built first; only the named source is recompiled with the extra flags)
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
NAME=$1; FLAGS=$2; SRC=${3:-conv_pfa108.cu}; BASE=${SRC%.cu}
OUT=$ROOT/build/variants; mkdir -p $OUT/obj_$NAME
cd $ROOT/fftconvolution_b200/csrc
ARCH="-gencode arch=compute_100a,code=sm_100a"
nvcc -O3 -std=c++17 -lineinfo $ARCH -Xcompiler -fPIC -Xptxas -v $FLAGS -I. -c $SRC -o $OUT/obj_$NAME/$BASE.o 2> $OUT/obj_$NAME/ptxas.log
grep -A2 "conv_kernel" $OUT/obj_$NAME/ptxas.log | grep -i "registers\|spill" | tr '\n' ' '; echo
OBJS=$(ls ../lib/obj/*.o | grep -v "/$BASE.o")
nvcc $ARCH -shared -o $OUT/$NAME.so $OBJS $OUT/obj_$NAME/$BASE.o -lcudart_static -lpthread -ldl -lrt
