"""Exploration: per-image kernels (kernel_is_shared = 0) at sizes whose work buffer fits shared memory. Not a test."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, fftconvolution_b200 as b2c
torch.cuda.set_device(0)
st = torch.cuda.current_stream().cuda_stream
PEAK = 6540.2e9
for (mode, hs, hk) in [(7, 20, 20), (7, 30, 30), (7, 50, 50), (2, 50, 10), (2, 100, 15), (7, 64, 64), (0, 100, 28), (2, 100, 100)]:
    n = 16384
    src = torch.empty((n, hs, hs), dtype=torch.float64, device="cuda"); b2c.fill_uniform_device(src.data_ptr(), src.numel(), 5)
    ker = torch.empty((n, hk, hk), dtype=torch.float64, device="cuda"); b2c.fill_uniform_device(ker.data_ptr(), ker.numel(), 6)
    p = b2c.Plan(mode, hs, hs, hk, hk)
    res = {}
    for path in (b2c.PATH_FFT, b2c.PATH_DIRECT):
        try:
            p.set_path(path)
            dst = torch.empty((n, p.h_dst, p.w_dst), dtype=torch.float64, device="cuda")
            for _ in range(2): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 0, dst.data_ptr(), n, st)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(3): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 0, dst.data_ptr(), n, st)
            e1.record(); torch.cuda.synchronize()
            res[path] = n / (e0.elapsed_time(e1) / 3) / 1e3
        except b2c.B2CError as e:
            res[path] = float("nan")
    b = 8 * (hs * hs + hk * hk + p.h_dst * p.w_dst)
    print("mode %d %dx%d (*) %dx%d per-image kernels P=%dx%d smem=%s: fft %.2f | direct %.2f M conv/s; best %.1f%% of roof" % (
        mode, hs, hs, hk, hk, p.h_fft, p.w_fft, p.info()["work_in_smem"], res[b2c.PATH_FFT], res[b2c.PATH_DIRECT], 100 * np.nanmax(list(res.values())) * 1e6 * b / PEAK), flush=True)
