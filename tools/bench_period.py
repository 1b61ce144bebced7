"""Throughput of the square tuned kernels at one period: python tools/bench_period.py P [P ...]
(LINEAR_FULL, source P-14, kernel 15, shared kernel; B2C_LIB_PATH selects a library variant)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, fftconvolution_b200 as b2c
torch.cuda.set_device(0)
for P in [int(a) for a in sys.argv[1:]]:
    hs = P - 14
    p = b2c.Plan(b2c.LINEAR_FULL, hs, hs, 15, 15); p.set_path(b2c.PATH_FFT)
    assert p.h_fft == P, (P, p.h_fft)
    n = int(max(4096, min(262144, 6e8 / (8 * (hs * hs + P * P)))))
    src = torch.rand((n, hs, hs), dtype=torch.float64, device="cuda"); ker = torch.rand((15, 15), dtype=torch.float64, device="cuda")
    dst = torch.empty((n, P, P), dtype=torch.float64, device="cuda")
    for _ in range(3): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, 0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, 0)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print("%s P=%d: %.2f ns/conv  %.2f M conv/s  checksum %x" % (os.path.basename(os.environ.get("B2C_LIB_PATH", "default")), P, ms * 1e6 / n, n / ms / 1e3, b2c.checksum_device(dst.data_ptr(), dst.numel())[0]))
    p.destroy()
