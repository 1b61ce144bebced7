#!/bin/bash
# 8-GPU node session: PCIe ceiling at 1/2/4/8 ranks, then bench.py at 8 (and 4) GPUs.
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O; : > $O/pcie_ceiling.jsonl
nvidia-smi topo -m > $O/topo.txt 2>&1
for N in 1 2 4 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N tools/pcie_ceiling.py >> $O/pcie_ceiling.jsonl 2>> $O/pcie.err
done
cat $O/pcie_ceiling.jsonl
for N in 8 4; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$N bench.py --gpus $N --steps 10 --warmup 3 > $O/bench_${N}gpu.json 2> $O/bench_${N}gpu.err
  cut -c1-1500 $O/bench_${N}gpu.json
done
