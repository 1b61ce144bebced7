#!/bin/bash
# GPU tests, then the direct-kernel table (extra environment, e.g. B2C_DIRECT_TILE=50, is passed through as a second run)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python tools/bench_direct.py 2>&1 | sed 's/^bulk/auto/' | tee gpurun_out/bench_direct.txt
for t in "$@"; do
  env $t timeout 300 python tools/bench_direct.py 2>&1 | sed "s/^bulk/$t/" | tee -a gpurun_out/bench_direct.txt
done
