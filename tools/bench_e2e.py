"""Host-buffer (pinned) end-to-end throughput of b2c_convolve_batch_host for the headline shape."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import fftconvolution_b200 as b2c
n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
torch.cuda.set_device(0)
h_src = torch.rand((n, 100, 100), dtype=torch.float64).pin_memory()
h_dst = torch.empty((n, 100, 100), dtype=torch.float64).pin_memory()
h_ker = torch.rand((15, 15), dtype=torch.float64).pin_memory()
plan = b2c.Plan(b2c.LINEAR_SAME, 100, 100, 15, 15)
plan.convolve_batch_host_ptr(h_src.data_ptr(), h_ker.data_ptr(), 1, h_dst.data_ptr(), n)
t0 = time.perf_counter()
for _ in range(3):
    plan.convolve_batch_host_ptr(h_src.data_ptr(), h_ker.data_ptr(), 1, h_dst.data_ptr(), n)
dt = (time.perf_counter() - t0) / 3
print("B2C_PIPE_MB=%s: %.1f ms  %.0f conv/s  %.1f GB/s each way" % (os.environ.get("B2C_PIPE_MB", "48"), dt * 1e3, n / dt, n * 80000 / dt / 1e9))
