"""BASELINE configs[3]/[4] slice for 1-D signals: FFT-vs-direct crossover on 1 x N rows (N x 1 columns run the same
kernels), LINEAR_SAME and LINEAR_FULL.  Columns as tools/crossover_all_modes.py without the host timings."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import fftconvolution_b200 as b2c
from bench import ClockSampler
torch.cuda.set_device(0)
st = torch.cuda.current_stream().cuda_stream

def rate(plan, path, src, ker, dst, n):
    plan.set_path(path)
    try:
        for _ in range(2): plan.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, st)
    except b2c.B2CError:
        return 0.0
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): plan.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, st)
    e1.record(); torch.cuda.synchronize()
    return n * 5 / (e0.elapsed_time(e1) * 1e-3)

sampler = ClockSampler(0); sampler.start()
print("mode,ns,nk,fft_conv_s,direct_conv_s,auto_conv_s,auto_path", flush=True)
for mode in (2, 0):
    for ns in (64, 256, 1000, 4096):
        for nk in (3, 5, 8, 12, 15, 16, 24, 31, 63):
            if nk > ns: continue
            plan = b2c.Plan(mode, 1, ns, 1, nk)
            n = int(max(2048, min(1 << 20, 3e8 / (8 * (ns + plan.w_dst)))))
            src = torch.empty((n, 1, ns), dtype=torch.float64, device="cuda"); b2c.fill_uniform_device(src.data_ptr(), src.numel(), 1)
            ker = torch.empty((1, nk), dtype=torch.float64, device="cuda"); b2c.fill_uniform_device(ker.data_ptr(), ker.numel(), 2)
            dst = torch.empty((n, 1, plan.w_dst), dtype=torch.float64, device="cuda")
            f = rate(plan, b2c.PATH_FFT, src, ker, dst, n); d = rate(plan, b2c.PATH_DIRECT, src, ker, dst, n)
            a = rate(plan, b2c.PATH_AUTO, src, ker, dst, n)
            ap = plan.info()["path"]
            plan.destroy()
            print("%d,%d,%d,%.4g,%.4g,%.4g,%d" % (mode, ns, nk, f, d, a, ap), flush=True)
            del src, ker, dst
sampler.sample(); sampler.stop_flag = True
print("# clocks %s" % (sampler.result(),))
