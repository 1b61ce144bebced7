#!/bin/bash
# parity tests, then the headline shape at several batch sizes (the kernel-spectrum launch shows in small batches)
cd "$(dirname "$0")/.."
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for n in 512 2500 8192 65536; do timeout 300 python tools/quick_bench.py $n 20 | sed "s/^/n=$n /"; done
