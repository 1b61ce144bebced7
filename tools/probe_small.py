"""Exploration: direct path on small images / short signals (config 5 corner, config 4 N=63/64). Not a test."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, fftconvolution_b200 as b2c
torch.cuda.set_device(0)
st = torch.cuda.current_stream().cuda_stream
PEAK = 6540.2e9
for (mode, hs, ws, hk, wk) in [(2, 100, 100, 1, 1), (2, 100, 100, 3, 3), (2, 100, 100, 5, 5), (2, 100, 100, 7, 7), (2, 70, 70, 3, 3), (2, 50, 50, 3, 3), (2, 50, 50, 5, 5), (2, 30, 30, 3, 3), (2, 20, 20, 7, 7), (2, 10, 10, 9, 9), (2, 10, 10, 3, 3), (4, 50, 50, 5, 5), (2, 1, 64, 1, 15), (2, 1, 4096, 1, 3), (2, 1, 4096, 1, 15)]:
    n = max(16384, min(1 << 20, (1 << 28) // (8 * hs * ws)))
    p = b2c.Plan(mode, hs, ws, hk, wk); p.set_path(b2c.PATH_DIRECT)
    src = torch.empty((n, hs, ws), dtype=torch.float64, device="cuda"); b2c.fill_uniform_device(src.data_ptr(), src.numel(), 5)
    ker = torch.rand((hk, wk), dtype=torch.float64, device="cuda")
    dst = torch.empty((n, p.h_dst, p.w_dst), dtype=torch.float64, device="cuda")
    for _ in range(2): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, st)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, st)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    b = 8 * (hs * ws + p.h_dst * p.w_dst)
    print("mode %d %dx%d (*) %dx%d n=%d: %.3f ms %.1f M conv/s  %.1f%% of HBM roof" % (mode, hs, ws, hk, wk, n, ms, n / ms / 1e3, 100 * n / ms * 1e3 * b / PEAK), flush=True)
