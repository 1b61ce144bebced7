#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python tools/crossover_all_modes.py > gpurun_out/crossover_all_modes.csv 2> gpurun_out/crossover.err
tail -5 gpurun_out/crossover_all_modes.csv; tail -3 gpurun_out/crossover.err
