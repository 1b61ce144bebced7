#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
B2C_P108=pfa timeout 300 python tools/trace_pfa108.py 2>&1 | tee gpurun_out/trace_pfa108_1cta.csv
