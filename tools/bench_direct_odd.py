"""Direct kernel on shapes the bulk-staged variant cannot take (odd widths, CIRCULAR_SAME wrapping):
python tools/bench_direct_odd.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, fftconvolution_b200 as b2c
torch.cuda.set_device(0)
tag = os.path.basename(os.environ.get("B2C_LIB_PATH", "default"))
for mode, hs, hk in [(2, 99, 3), (2, 99, 5), (2, 99, 7), (2, 63, 3), (2, 49, 3), (2, 49, 7), (2, 29, 3), (2, 199, 3), (4, 100, 3), (4, 100, 5), (4, 50, 3)]:
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
    print("%s: %s %dx%d (*) %dx%d: %.2f M conv/s  %.0f GB/s = %.1f%% of the measured HBM roof" % (tag, b2c.MODE_NAMES[mode], hs, hs, hk, hk, n / ms / 1e3, gbs, gbs / 65.306))
    p.destroy()
