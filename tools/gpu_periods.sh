#!/bin/bash
# CTA-size A/B of the square tuned kernels: default library on all periods, then each build/variants/tt_P_T.so on its period
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
{ timeout 600 python tools/bench_period.py ${PERIODS:-16 20 24 28 32 36 40 42 48 50 54 56 60 64 70 72 80 84 90 96 100 108 112 120 126 128 140 144 160}
  for f in $(ls build/variants/tt_*.so 2>/dev/null); do
    P=$(basename $f .so | cut -d_ -f2)
    B2C_LIB_PATH=$f timeout 300 python tools/bench_period.py $P
  done; } 2>&1 | tee gpurun_out/periods.txt
