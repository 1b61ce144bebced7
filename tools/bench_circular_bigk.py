"""CIRCULAR_SAME with kernels larger than the source (they wrap around it): python tools/bench_circular_bigk.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, fftconvolution_b200 as b2c
torch.cuda.set_device(0)
for s, k in [(43, 61), (61, 70), (61, 100), (80, 100), (97, 99), (97, 80), (89, 85), (30, 100)]:
    p = b2c.Plan(b2c.CIRCULAR_SAME, s, s, k, k); p.set_path(b2c.PATH_FFT)
    n = 4096
    src = torch.rand((n, s, s), dtype=torch.float64, device="cuda"); ker = torch.rand((k, k), dtype=torch.float64, device="cuda")
    dst = torch.empty((n, s, s), dtype=torch.float64, device="cuda")
    for _ in range(2): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, 0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, 0)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print("CIRCULAR_SAME %dx%d (*) %dx%d shared: %.1f ns/conv  %.2f M conv/s" % (s, s, k, k, ms * 1e6 / n, n / ms / 1e3))
    p.destroy()
