#!/bin/bash
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
CMD="timeout 300 python tools/run_direct_once.py ${1:-3} ${2:-100}"
$CMD > $O/plain_direct.log 2>&1 && cat $O/plain_direct.log &&
ncu --set full --clock-control none --import-source on -k regex:direct_ -s 2 -c 1 -o $O/ncu_direct -f $CMD > $O/ncu_direct.log 2>&1
ls -la $O | grep ncu_direct
