#!/bin/bash
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
CMD="timeout 300 python tools/bench_config3.py"
$CMD > $O/plain_c3.log 2>&1 && cat $O/plain_c3.log &&
ncu --set full --clock-control none --import-source on -k regex:conv_kernel -s 2 -c 1 -o $O/ncu_c3 -f $CMD > $O/ncu_c3.log 2>&1
ls -la $O | grep ncu_c3
