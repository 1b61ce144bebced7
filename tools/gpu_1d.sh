#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out; : > gpurun_out/bench_1d.txt
for v in "$@"; do
  if [ "$v" = default ]; then timeout 300 python tools/bench_1d.py >> gpurun_out/bench_1d.txt 2>&1
  else B2C_LIB_PATH=build/variants/$v.so timeout 300 python tools/bench_1d.py >> gpurun_out/bench_1d.txt 2>&1; fi
done
cat gpurun_out/bench_1d.txt
