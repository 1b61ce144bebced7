"""Quick kernel-only A/B timing of the headline workload (not the bench contract; see bench.py).
usage: [B2C_LIB_PATH=variant.so] python tools/quick_bench.py [n_images] [steps]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import fftconvolution_b200 as b2c

n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
torch.cuda.set_device(0)
src = torch.empty((n, 100, 100), dtype=torch.float64, device="cuda")
dst = torch.empty_like(src)
b2c.fill_uniform_device(src.data_ptr(), src.numel(), 0x5EED0002)
ker = torch.from_numpy(np.random.default_rng(0).random((15, 15))).cuda()
plan = b2c.Plan(b2c.LINEAR_SAME, 100, 100, 15, 15)
st = torch.cuda.current_stream().cuda_stream
for _ in range(3):
    plan.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, st)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(steps):
    plan.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, st)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / steps
cs = b2c.checksum_device(dst.data_ptr(), dst.numel())
print("%s: %.3f ms/step  %.2f M conv/s  checksum %x" % (os.environ.get("B2C_LIB_PATH", "default"), ms, n / ms / 1e3, cs[0]))
