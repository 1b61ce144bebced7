#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
B2C_DIRECT_NBUF=1 timeout 300 python tools/bench_direct.py 2>&1 | sed 's/^bulk/nbuf1/' | tee gpurun_out/bench_direct.txt
B2C_DIRECT_NBUF=2 timeout 300 python tools/bench_direct.py 2>&1 | sed 's/^bulk/nbuf2/' | tee -a gpurun_out/bench_direct.txt
timeout 300 python tools/bench_direct.py 2>&1 | sed 's/^bulk/auto/' | tee -a gpurun_out/bench_direct.txt
