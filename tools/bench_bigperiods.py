"""Throughput of LINEAR_FULL 100 x 100 (*) k x k for kernels whose period lies beyond the shared-memory kernels:
python tools/bench_bigperiods.py [k ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, fftconvolution_b200 as b2c
torch.cuda.set_device(0)
ks = [int(a) for a in sys.argv[1:]] or [45, 61, 65, 70, 80, 90, 95, 100]
for k in ks:
    for shared in (1, 0):
        p = b2c.Plan(b2c.LINEAR_FULL, 100, 100, k, k); p.set_path(b2c.PATH_FFT)
        n = 2048
        src = torch.rand((n, 100, 100), dtype=torch.float64, device="cuda")
        ker = torch.rand((k, k) if shared else (n, k, k), dtype=torch.float64, device="cuda")
        dst = torch.empty((n, p.h_dst, p.w_dst), dtype=torch.float64, device="cuda")
        for _ in range(2): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), shared, dst.data_ptr(), n, 0)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), shared, dst.data_ptr(), n, 0)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        print("100x100 (*) %dx%d LINEAR_FULL %s: period %dx%d  %.1f ns/conv  %.2f M conv/s  work_in_smem %d" % (k, k, "shared" if shared else "per-image", p.h_fft, p.w_fft, ms * 1e6 / n, n / ms / 1e3, p.info()["work_in_smem"]))
        p.destroy()
