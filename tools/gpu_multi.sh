#!/bin/bash
# N-GPU session: the parity test that needs more than one device, then bench.py under torchrun at N (and the reference arm).
cd "$(dirname "$0")/.."
N=${1:-2}
O=gpurun_out; mkdir -p $O
nvidia-smi -L > $O/smi_multi.txt 2>&1
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "multi_device" > $O/pytest_multi.log 2>&1; tail -3 $O/pytest_multi.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > $O/bench_${N}gpu.json 2> $O/bench_${N}gpu.err
cat $O/bench_${N}gpu.json | cut -c1-3000; tail -3 $O/bench_${N}gpu.err
