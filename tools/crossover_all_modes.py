"""BASELINE configs[4] slice: FFT-vs-direct crossover for ALL eight modes on square shapes, with the reference's own
CPU paths timed on the host cores for the same points (reference grid: run_fftw_bench.sh:13-19).
Columns: mode,hs,hk,fft,direct,auto (conv/s on the GPU, batch resident in HBM), auto_path (1 fft / 2 direct),
host_fft (unmodified convolution_fftw.h over the stand-in FFT, all cores), host_loops (unmodified convolution_std.h,
all cores; n/a for modes it does not have).  The last lines record clocks and core count."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import fftconvolution_b200 as b2c
import _oracle
from bench import ClockSampler, host_cores

torch.cuda.set_device(0)
st = torch.cuda.current_stream().cuda_stream
ref = _oracle.load_ref()
cores = host_cores()
STD_MODE = {0: 0, 1: 1, 2: 1, 3: 2, 4: 3, 6: 4, 7: 4}   # FFTW_Convolution mode -> STD_Convolution mode (convolution_std.h:10-17)


def rate(plan, path, src, ker, dst, n):
    plan.set_path(path)
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


sampler = ClockSampler(0); sampler.start()
print("mode,hs,hk,fft_conv_s,direct_conv_s,auto_conv_s,auto_path,host_fft_conv_s,host_loops_conv_s", flush=True)
rng = np.random.default_rng(5)
for mode in range(8):
    for s in (10, 30, 50, 100):
        for k in (1, 3, 5, 7, 9, 11, 15, 21, 31):
            plan = b2c.Plan(mode, s, s, k, k)
            if plan.h_dst == 0:
                plan.destroy(); continue
            n = int(max(512, min(262144, 3e8 / (8 * (s * s + plan.h_dst * plan.w_dst)))))
            src = torch.empty((n, s, s), dtype=torch.float64, device="cuda"); b2c.fill_uniform_device(src.data_ptr(), src.numel(), 1)
            ker = torch.empty((k, k), dtype=torch.float64, device="cuda"); b2c.fill_uniform_device(ker.data_ptr(), ker.numel(), 2)
            dst = torch.empty((n, plan.h_dst, plan.w_dst), dtype=torch.float64, device="cuda")
            f = rate(plan, b2c.PATH_FFT, src, ker, dst, n)
            d = rate(plan, b2c.PATH_DIRECT, src, ker, dst, n)
            a = rate(plan, b2c.PATH_AUTO, src, ker, dst, n)
            ap = plan.info()["path"]
            plan.destroy()
            hf = hl = float("nan")
            if ref is not None:
                nh = int(max(4 * cores, min(2048, 2e6 / (s * s * max(k, 4)))))
                hs_, hk_ = rng.random((nh, s, s)), rng.random((k, k))
                t0 = time.perf_counter(); ref.batch_fftw(mode, hs_, hk_, threads=cores); hf = nh / (time.perf_counter() - t0)
                if mode in STD_MODE:
                    t0 = time.perf_counter(); ref.batch_std(STD_MODE[mode], hs_, hk_, threads=cores); hl = nh / (time.perf_counter() - t0)
            print("%d,%d,%d,%.4g,%.4g,%.4g,%d,%.4g,%.4g" % (mode, s, k, f, d, a, ap, hf, hl), flush=True)
            del src, ker, dst
sampler.sample(); sampler.stop_flag = True
print("# clocks %s; host cores %d; host paths: reference headers unmodified, FFT header over the stand-in FFT (no libfftw3 on the box)" % (sampler.result(), cores))
