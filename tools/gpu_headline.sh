#!/bin/bash
# One GPU session for the headline kernel: parity tests, A/B bench against the seven-pass kernel, per-pass clock
# trace, bench line, then (only if everything before exited 0) the ncu launch list and one full ncu capture.
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $O/smi.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; rc=$?; echo "pytest exit $rc" >> $O/pytest_gpu.log
tail -5 $O/pytest_gpu.log
timeout 300 python tools/quick_bench.py 65536 20 > $O/ab.txt 2>&1
B2C_P108=pfa timeout 300 python tools/quick_bench.py 65536 20 >> $O/ab.txt 2>&1
cat $O/ab.txt
[ $rc -eq 0 ] || { echo "tests failed: no profiling"; exit 1; }
true
timeout 600 python bench.py --steps 20 --warmup 3 > $O/bench.json 2> $O/bench.err; cat $O/bench.json
CMD="timeout 300 python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-secondary"
$CMD > $O/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/launches.csv $CMD > $O/ncu_launch.log 2>&1
$CMD > $O/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:conv_kernel -s 3 -c 1 -o $O/pfa108_full -f $CMD > $O/ncu_full.log 2>&1
ls -la $O
