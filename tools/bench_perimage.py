"""Per-image-kernel throughput of square shapes: python tools/bench_perimage.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, fftconvolution_b200 as b2c
torch.cuda.set_device(0)
for hs, hk, n in [(100, 15, 32768), (50, 10, 65536), (100, 5, 32768), (64, 31, 32768)]:
    p = b2c.Plan(b2c.LINEAR_SAME, hs, hs, hk, hk)
    src = torch.rand((n, hs, hs), dtype=torch.float64, device="cuda"); ker = torch.rand((n, hk, hk), dtype=torch.float64, device="cuda")
    dst = torch.empty_like(src)
    for path in (b2c.PATH_AUTO, b2c.PATH_FFT, b2c.PATH_FFT_GENERIC):
        p.set_path(path)
        for _ in range(3): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 0, dst.data_ptr(), n, 0)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 0, dst.data_ptr(), n, 0)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print("%dx%d (*) %dx%d per-image kernels, path %d (period %d): %.3f ms  %.2f M conv/s" % (hs, hs, hk, hk, path, p.h_fft, ms, n / ms / 1e3))
    p.destroy()
