#!/bin/bash
# ncu full capture of the headline kernel of one library variant: tools/gpu_ncu_variant.sh NAME
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
export B2C_LIB_PATH=build/variants/$1.so
CMD="timeout 300 python tools/quick_bench.py 65536 3"
$CMD > $O/plain_$1.log 2>&1 && cat $O/plain_$1.log &&
ncu --set full --clock-control none --import-source on -k regex:conv_kernel -s 3 -c 1 -o $O/ncu_$1 -f $CMD > $O/ncu_$1.log 2>&1
ls -la $O | grep $1
