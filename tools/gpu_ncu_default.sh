#!/bin/bash
# ncu full capture (with source) of the headline kernel as the in-tree library runs it
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
CMD="timeout 300 python tools/quick_bench.py 65536 3"
$CMD > $O/plain_default.log 2>&1 && cat $O/plain_default.log &&
ncu --set full --clock-control none --import-source on -k regex:conv_kernel -s 3 -c 1 -o $O/ncu_default -f $CMD > $O/ncu_default.log 2>&1
ls -la $O | grep ncu_default
