#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python tools/crossover_all_modes.py > gpurun_out/crossover_all_modes.csv 2> gpurun_out/crossover.err
tail -2 gpurun_out/crossover_all_modes.csv
timeout 1500 python tools/emit_benchmark_table.py 100 1 > gpurun_out/benchmarks_convolution_b2c.txt 2> gpurun_out/table.err
wc -l gpurun_out/benchmarks_convolution_b2c.txt
