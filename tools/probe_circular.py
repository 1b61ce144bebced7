"""Exploration: CIRCULAR_SAME, tuned menu period (periodic extension) vs the any-size kernel. Not a test."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, fftconvolution_b200 as b2c
torch.cuda.set_device(0)
st = torch.cuda.current_stream().cuda_stream
PEAK = 6540.2e9
for (hs, ws, hk, wk) in [(31, 37, 5, 7), (50, 50, 5, 5), (63, 63, 9, 9), (97, 89, 9, 3), (100, 100, 9, 9), (100, 100, 15, 15), (1, 1000, 1, 255), (1, 997, 1, 64), (1, 4000, 1, 63)]:
    n = max(4096, min(262144, (1 << 27) // (8 * hs * ws)))
    src = torch.empty((n, hs, ws), dtype=torch.float64, device="cuda"); b2c.fill_uniform_device(src.data_ptr(), src.numel(), 5)
    ker = torch.rand((hk, wk), dtype=torch.float64, device="cuda")
    p = b2c.Plan(b2c.CIRCULAR_SAME, hs, ws, hk, wk)
    res = []
    for path in (b2c.PATH_FFT, b2c.PATH_FFT_GENERIC, b2c.PATH_DIRECT):
        try:
            p.set_path(path)
            dst = torch.empty((n, hs, ws), dtype=torch.float64, device="cuda")
            for _ in range(2): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, st)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, st)
            e1.record(); torch.cuda.synchronize()
            res.append(n / (e0.elapsed_time(e1) / 5) / 1e3)
        except b2c.B2CError:
            res.append(float("nan"))
    p.set_path(b2c.PATH_AUTO)
    print("%dx%d (*) %dx%d own period %dx%d: tuned %.1f | any-size %.1f | direct %.1f M conv/s; auto path %d; best %.1f%% of roof" % (
        hs, ws, hk, wk, p.h_fft, p.w_fft, res[0], res[1], res[2], p.info()["path"], 100 * np.nanmax(res) * 1e6 * 16 * hs * ws / PEAK), flush=True)
