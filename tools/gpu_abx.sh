#!/bin/bash
# A/B of library variants on one bench script: tools/gpu_abx.sh tools/bench_direct.py default d_ti10 ...
# (environment such as B2C_DIRECT_CW=1 is passed through)
cd "$(dirname "$0")/.."
S=$1; shift
mkdir -p gpurun_out; : > gpurun_out/abx.txt
for v in "$@"; do
  echo "== $v" >> gpurun_out/abx.txt
  if [ "$v" = default ]; then timeout 300 python $S >> gpurun_out/abx.txt 2>&1
  else B2C_LIB_PATH=build/variants/$v.so timeout 300 python $S >> gpurun_out/abx.txt 2>&1; fi
done
cat gpurun_out/abx.txt
