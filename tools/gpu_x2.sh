#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out; : > gpurun_out/ab.txt
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "tuned_108" 2>&1 | tail -3
for k in x2 t2 pfa; do B2C_P108=$k timeout 300 python tools/quick_bench.py 65536 20 2>&1 | sed "s/^default/$k/" | tee -a gpurun_out/ab.txt; done
