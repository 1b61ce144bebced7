#!/bin/bash
# config 3 (100x100 (*) 100x100 CIRCULAR_FULL, per-image kernels): strip-mode parity tests and throughput, per library variant
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for v in "$@"; do
  if [ "$v" != default ]; then export B2C_LIB_PATH=build/variants/$v.so; else unset B2C_LIB_PATH; fi
  echo "$v $(timeout 900 python -m pytest tests -m gpu -x -q -k 'config3 or strip or packed or fuzz' 2>&1 | tail -1)"
  timeout 300 python tools/bench_config3.py 2>&1 | sed "s/^/$v /"
done | tee gpurun_out/bench_config3.txt
