#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "golden or edge_sizes or kernel_cache or harness or one_d or categories" 2>&1 | tail -3
echo "--- copies"; B2C_ZEROCOPY=0 timeout 300 python tools/probe_latency.py 2>&1 | grep -E "path 0|path 1:|path 2:" | tee gpurun_out/latency_copies.txt
echo "--- zero copy"; timeout 300 python tools/probe_latency.py 2>&1 | grep -E "path 0|path 1:|path 2:" | tee gpurun_out/latency_zerocopy.txt
