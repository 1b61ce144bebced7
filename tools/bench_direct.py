"""Direct-kernel throughput on square shapes: python tools/bench_direct.py   (B2C_DIRECT_NO_BULK=1 for the element-wise staging)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, fftconvolution_b200 as b2c
torch.cuda.set_device(0)
tag = "nobulk" if os.environ.get("B2C_DIRECT_NO_BULK") else "bulk"
for mode, hs, hk in [(2, 100, 1), (2, 100, 3), (2, 100, 5), (2, 100, 7), (2, 100, 9), (0, 100, 3), (3, 100, 5), (2, 64, 3), (2, 50, 3), (2, 50, 7), (2, 30, 3), (2, 200, 3)]:
    p = b2c.Plan(mode, hs, hs, hk, hk); p.set_path(b2c.PATH_DIRECT)
    n = int(max(2048, min(262144, 4e8 / (8 * (hs * hs + p.h_dst * p.w_dst)))))
    src = torch.rand((n, hs, hs), dtype=torch.float64, device="cuda"); ker = torch.rand((hk, hk), dtype=torch.float64, device="cuda")
    dst = torch.empty((n, p.h_dst, p.w_dst), dtype=torch.float64, device="cuda")
    for _ in range(3): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, 0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, 0)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    gbs = n * 8 * (hs * hs + p.h_dst * p.w_dst) / ms / 1e6
    print("%s: mode %d %dx%d (*) %dx%d: %.3f ms  %.2f M conv/s  %.0f GB/s = %.1f%% of the measured HBM roof" % (tag, mode, hs, hs, hk, hk, ms, n / ms / 1e3, gbs, gbs / 65.306))
    p.destroy()
