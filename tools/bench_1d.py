"""1-D throughput (BASELINE configs[3] shapes): python tools/bench_1d.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, fftconvolution_b200 as b2c
torch.cuda.set_device(0)
for mode, ns, nk, n in [(2, 4096, 255, 16384), (0, 4095, 256, 16384), (2, 1000, 255, 65536), (2, 4096, 4096, 16384), (2, 100, 40, 262144)]:
    p = b2c.Plan(mode, 1, ns, 1, nk)
    src = torch.rand((n, 1, ns), dtype=torch.float64, device="cuda"); ker = torch.rand((1, nk), dtype=torch.float64, device="cuda")
    dst = torch.empty((n, p.h_dst, p.w_dst), dtype=torch.float64, device="cuda")
    for _ in range(3): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, 0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, 0)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    gbs = n * 8 * (ns + p.w_dst) / ms / 1e6
    print("%s: mode %d 1x%d (*) 1x%d period %d: %.3f ms  %.2f M conv/s  %.0f GB/s (%.1f%% of 6530)" % (os.environ.get("B2C_LIB_PATH", "default"), mode, ns, nk, p.w_fft, ms, n / ms / 1e3, gbs, gbs / 65.306))
    p.destroy()
