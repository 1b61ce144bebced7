#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for i in 1 2 3; do timeout 300 python tools/quick_bench.py 65536 20 2>&1 | tee -a gpurun_out/ab.txt; done
