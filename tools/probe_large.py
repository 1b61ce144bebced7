"""Exploration: large-image sizes through the strip-mode path. Not a test."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, fftconvolution_b200 as b2c
from scipy.signal import fftconvolve
torch.cuda.set_device(0)
rng = np.random.default_rng(5)
for (hs, hk, n) in [(4096, 15, 1), (4096, 63, 2), (4500, 129, 1)]:
    try:
        p = b2c.Plan(b2c.LINEAR_SAME, hs, hs, hk, hk); p.set_path(b2c.PATH_FFT)
        info = p.info()
        src = rng.random((n, hs, hs)); ker = rng.random((hk, hk))
        d_src = torch.from_numpy(src).cuda(); d_ker = torch.from_numpy(ker).cuda()
        d_dst = torch.empty((n, hs, hs), dtype=torch.float64, device="cuda")
        p.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), 1, d_dst.data_ptr(), n, 0); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); p.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), 1, d_dst.data_ptr(), n, 0); e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        got = d_dst[n - 1].cpu().numpy()
        full = fftconvolve(src[n - 1], ker, mode="full")
        ref = full[hk // 2: hk // 2 + hs, hk // 2: hk // 2 + hs]
        err = np.abs(got - ref).max() / np.abs(ref).max()
        gbs = n * 2 * hs * hs * 8 / ms / 1e6
        print(f"{hs}x{hs} (*) {hk}x{hk} n={n}: P={p.h_fft}x{p.w_fft} path={info.get('path')} {ms:.2f} ms  {n/ms*1e3:.0f} conv/s  {gbs:.0f} GB/s algorithmic  relerr {err:.1e}", flush=True)
        p.destroy()
    except Exception as e:
        print(f"{hs} (*) {hk}: FAILED {type(e).__name__}: {e}", flush=True)
