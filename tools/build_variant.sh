#!/bin/bash
# Build an A/B variant of libb2conv.so: tools/build_variant.sh NAME "-DFLAG=1 ..."  -> build/variants/NAME.so
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
NAME=$1; FLAGS=$2; SRC=${3:-conv_tuned108.cu}
OUT=$ROOT/build/variants; mkdir -p $OUT/obj_$NAME
cd $ROOT/fftconvolution_b200/csrc
ARCH="-gencode arch=compute_100a,code=sm_100a"
nvcc -O3 -std=c++17 -lineinfo $ARCH -Xcompiler -fPIC -Xptxas -v $FLAGS -I. -c $SRC -o $OUT/obj_$NAME/conv_tuned108.o 2> $OUT/obj_$NAME/ptxas.log
grep -A2 "conv_kernel" $OUT/obj_$NAME/ptxas.log | grep -i "registers\|spill" | tr '\n' ' '; echo
nvcc $ARCH -shared -o $OUT/$NAME.so ../lib/obj/capi.o ../lib/obj/conv_generic.o ../lib/obj/conv_direct.o ../lib/obj/conv_big2d.o ../lib/obj/conv_big2ds.o $OUT/obj_$NAME/conv_tuned108.o ../lib/obj/conv_tuned2.o ../lib/obj/conv_tuned1d.o ../lib/obj/static_scratch_entry.o ../lib/obj/util_kernels.o ../lib/obj/planner.o -lcudart_static -lpthread -ldl -lrt
