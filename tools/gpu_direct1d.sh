#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
B2C_DIRECT_NO_BULK=1 timeout 300 python tools/bench_direct1d.py 2>&1 | tee gpurun_out/bench_direct1d.txt
timeout 300 python tools/bench_direct1d.py 2>&1 | tee -a gpurun_out/bench_direct1d.txt
