"""Emit GPU timings in the format of the reference's published tables (Convolution/Data/benchmarks_convolution_*.txt:
`mode src kernel seconds`, src 4..MAX, kernel 3..src-1, square, modes linear / circular / linear_unpadded /
circular_unpadded as in convolution_benchmark.cc:70-97) so the reference's plot scripts can overlay them.
`seconds` is the batch-amortised time of ONE convolution on one B200 (device-resident batch).
usage: python tools/emit_benchmark_table.py [MAX=100] [STEP=1] > benchmarks_convolution_b2c.txt"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import fftconvolution_b200 as b2c

MAX = int(sys.argv[1]) if len(sys.argv) > 1 else 100
STEP = int(sys.argv[2]) if len(sys.argv) > 2 else 1
MODES = [("linear", b2c.LINEAR_SAME), ("circular", b2c.CIRCULAR_FULL), ("linear_unpadded", b2c.LINEAR_SAME_UNPADDED),
         ("circular_unpadded", b2c.CIRCULAR_FULL_UNPADDED)]
torch.cuda.set_device(0)
st = torch.cuda.current_stream().cuda_stream
big = torch.empty(64 << 20, dtype=torch.float64, device="cuda")      # 512 MB arena: src | dst
b2c.fill_uniform_device(big.data_ptr(), big.numel(), 42)
ker = torch.empty(100 * 100, dtype=torch.float64, device="cuda")
b2c.fill_uniform_device(ker.data_ptr(), ker.numel(), 43)
for name, mode in MODES:
    for s in range(4, MAX + 1, STEP):
        for k in range(3, s, STEP):
            plan = b2c.Plan(mode, s, s, k, k)
            per = s * s + plan.h_dst * plan.w_dst
            n = int(max(256, min(65536, (32 << 20) // per)))
            src_ptr, dst_ptr = big.data_ptr(), big.data_ptr() + 8 * n * s * s
            plan.convolve_batch_device(src_ptr, ker.data_ptr(), 1, dst_ptr, n, st)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            plan.convolve_batch_device(src_ptr, ker.data_ptr(), 1, dst_ptr, n, st)
            plan.convolve_batch_device(src_ptr, ker.data_ptr(), 1, dst_ptr, n, st)
            e1.record(); torch.cuda.synchronize()
            print("%s %d %d %e" % (name, s, k, e0.elapsed_time(e1) * 1e-3 / (2 * n)), flush=True)
            plan.destroy()
