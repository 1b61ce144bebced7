"""Measured FFT-vs-direct crossover (BASELINE config 5 slice + 1-D).  Prints CSV: kind,hs,ws,hk,wk,fft,direct"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import fftconvolution_b200 as b2c
torch.cuda.set_device(0)
st = torch.cuda.current_stream().cuda_stream

def rate(mode, hs, ws, hk, wk, path, target_bytes=4e8):
    plan = b2c.Plan(mode, hs, ws, hk, wk); plan.set_path(path)
    n = int(max(512, min(262144, target_bytes / (8 * (hs * ws + plan.h_dst * plan.w_dst)))))
    src = torch.empty((n, hs, ws), dtype=torch.float64, device="cuda"); b2c.fill_uniform_device(src.data_ptr(), src.numel(), 1)
    ker = torch.empty((hk, wk), dtype=torch.float64, device="cuda"); b2c.fill_uniform_device(ker.data_ptr(), ker.numel(), 2)
    dst = torch.empty((n, plan.h_dst, plan.w_dst), dtype=torch.float64, device="cuda")
    try:
        for _ in range(2): plan.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, st)
    except b2c.B2CError:
        return 0.0
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): plan.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, st)
    e1.record(); torch.cuda.synchronize()
    return n * 3 / (e0.elapsed_time(e1) * 1e-3)

print("kind,hs,ws,hk,wk,fft_conv_s,direct_conv_s,winner")
for s in (10, 20, 30, 50, 70, 100):
    for k in (1, 3, 5, 7, 9, 11, 13, 15, 17, 21):
        if k > s: continue
        f = rate(b2c.LINEAR_SAME, s, s, k, k, b2c.PATH_FFT); d = rate(b2c.LINEAR_SAME, s, s, k, k, b2c.PATH_DIRECT)
        print("2d,%d,%d,%d,%d,%.4g,%.4g,%s" % (s, s, k, k, f, d, "direct" if d > f else "fft"), flush=True)
for n in (64, 1000, 4096):
    for m in (3, 8, 15, 16, 32, 64, 128, 256):
        if m > n: continue
        f = rate(b2c.LINEAR_SAME, 1, n, 1, m, b2c.PATH_FFT); d = rate(b2c.LINEAR_SAME, 1, n, 1, m, b2c.PATH_DIRECT)
        print("1d_row,1,%d,1,%d,%.4g,%.4g,%s" % (n, m, f, d, "direct" if d > f else "fft"), flush=True)
        f = rate(b2c.LINEAR_SAME, n, 1, m, 1, b2c.PATH_FFT); d = rate(b2c.LINEAR_SAME, n, 1, m, 1, b2c.PATH_DIRECT)
        print("1d_col,%d,1,%d,1,%.4g,%.4g,%s" % (n, m, f, d, "direct" if d > f else "fft"), flush=True)
