#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python tools/crossover_1d.py > gpurun_out/crossover_1d.csv 2> gpurun_out/crossover_1d.err; tail -3 gpurun_out/crossover_1d.csv
