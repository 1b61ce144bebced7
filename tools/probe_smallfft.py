"""Exploration: FFT path on small square periods (tuned family vs any-size kernel). Not a test."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, fftconvolution_b200 as b2c
torch.cuda.set_device(0)
st = torch.cuda.current_stream().cuda_stream
PEAK = 6540.2e9
for (hs, hk) in [(36, 17), (50, 10), (56, 15), (64, 15), (70, 15), (80, 15), (86, 15), (90, 15), (100, 21), (110, 15), (118, 15)]:
    n = 131072
    src = torch.empty((n, hs, hs), dtype=torch.float64, device="cuda"); b2c.fill_uniform_device(src.data_ptr(), src.numel(), 5)
    ker = torch.rand((hk, hk), dtype=torch.float64, device="cuda")
    res = []
    for path in (b2c.PATH_FFT, b2c.PATH_FFT_GENERIC, b2c.PATH_DIRECT):
        p = b2c.Plan(2, hs, hs, hk, hk); p.set_path(path)
        dst = torch.empty((n, hs, hs), dtype=torch.float64, device="cuda")
        try:
            for _ in range(2): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, st)
        except b2c.B2CError:
            res.append(float("nan")); continue
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, st)
        e1.record(); torch.cuda.synchronize()
        res.append(n / (e0.elapsed_time(e1) / 5) / 1e3)
        pf = p.h_fft
    b = 16 * hs * hs
    print("%dx%d (*) %dx%d P=%d: tuned %.1f | any-size %.1f | direct %.1f M conv/s; best %.1f%% of roof" % (hs, hs, hk, hk, pf, res[0], res[1], res[2], 100 * np.nanmax(res) * 1e6 * b / PEAK), flush=True)
