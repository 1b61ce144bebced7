"""Measure BASELINE.json configs 1, 3, 4, 5 (config 2 is bench.py's headline).  Prints one JSON line per
point; rooflines: algorithmic bytes per convolution of SURVEY.md 8d over the measured HBM copy bandwidth.
usage: python tools/bench_configs.py [--quick]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import fftconvolution_b200 as b2c

QUICK = "--quick" in sys.argv
torch.cuda.set_device(0)
try:
    PEAK = json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"]
except Exception:
    PEAK = 6650.0
st = torch.cuda.current_stream().cuda_stream


def time_batch(mode, hs, ws, hk, wk, n, shared, path=b2c.PATH_AUTO, steps=5):
    plan = b2c.Plan(mode, hs, ws, hk, wk)
    plan.set_path(path)
    if plan.h_dst == 0:
        return None
    src = torch.empty((n, hs, ws), dtype=torch.float64, device="cuda")
    b2c.fill_uniform_device(src.data_ptr(), src.numel(), 0x5EED0003)
    ker = torch.empty((hk, wk) if shared else (n, hk, wk), dtype=torch.float64, device="cuda")
    b2c.fill_uniform_device(ker.data_ptr(), ker.numel(), 0x5EED0004)
    dst = torch.empty((n, plan.h_dst, plan.w_dst), dtype=torch.float64, device="cuda")
    try:
        for _ in range(2):
            plan.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1 if shared else 0, dst.data_ptr(), n, st)
    except b2c.B2CError as e:
        return {"error": str(e)[:80]}
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        plan.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1 if shared else 0, dst.data_ptr(), n, st)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    bytes_per = 8 * (hs * ws + plan.h_dst * plan.w_dst + (0 if shared else hk * wk))
    rate = n / (ms * 1e-3)
    info = plan.info()
    return {"conv_per_s": rate, "ms": ms, "n": n, "bytes_per_conv": bytes_per, "GBps": rate * bytes_per / 1e9,
            "hbm_frac": rate * bytes_per / 1e9 / PEAK, "fft": [plan.h_fft, plan.w_fft], "path": info["path"],
            "work_in_smem": info["work_in_smem"]}


def emit(tag, d):
    print(json.dumps({"config": tag, **(d or {"skipped": True})}), flush=True)


# config 1: single 50x50 (*) 10x10 LINEAR_SAME through the host entry point (latency), rand()-style inputs
rng = np.random.default_rng(1)
src, ker = rng.random((50, 50)), rng.random((10, 10))
plan = b2c.Plan(b2c.LINEAR_SAME, 50, 50, 10, 10)
for _ in range(5):
    plan.convolve_host(src, ker)
t0 = time.perf_counter()
for _ in range(200):
    plan.convolve_host(src, ker)
lat = (time.perf_counter() - t0) / 200
emit("1: single 50x50(*)10x10 LINEAR_SAME via convolve() host path", {"seconds_per_call": lat, "path": plan.info()["path"]})
emit("1b: same shape, device-resident batch 65536", time_batch(b2c.LINEAR_SAME, 50, 50, 10, 10, 65536, True))

# config 3: per-image kernels 100x100 (*) 100x100 CIRCULAR_FULL
emit("3: 100x100(*)100x100 CIRCULAR_FULL per-image kernels", time_batch(b2c.CIRCULAR_FULL, 100, 100, 100, 100, 2048 if QUICK else 16384, False, steps=3))

# config 4: 1-D
for n1 in ([4096] if QUICK else [63, 64, 1000, 4095, 4096]):
    for m in ([15, 256] if QUICK else [3, 4, 15, 16, 255, 256, n1]):
        for mode in (b2c.LINEAR_SAME, b2c.LINEAR_FULL, b2c.CIRCULAR_SAME):
            emit("4: 1x%d (*) 1x%d %s shared" % (n1, m, b2c.MODE_NAMES[mode]), time_batch(mode, 1, n1, 1, m, 16384, True, steps=3))

# config 5 (slice): source 100x100, kernel k x k, LINEAR_SAME: FFT vs direct crossover
for k in ([3, 9, 15] if QUICK else [1, 3, 5, 7, 9, 11, 13, 15, 17, 21, 25, 31]):
    f = time_batch(b2c.LINEAR_SAME, 100, 100, k, k, 16384, True, b2c.PATH_FFT, steps=3)
    d = time_batch(b2c.LINEAR_SAME, 100, 100, k, k, 16384, True, b2c.PATH_DIRECT, steps=3)
    emit("5: 100x100 (*) %dx%d LINEAR_SAME" % (k, k), {"fft_conv_per_s": f and f.get("conv_per_s"), "direct_conv_per_s": d and d.get("conv_per_s"),
                                                     "fft_period": f and f.get("fft"), "auto_path": b2c.Plan(b2c.LINEAR_SAME, 100, 100, k, k).info()["path"]})
for s in ([30] if QUICK else [10, 30, 50]):
    for k in ([3, 9] if QUICK else [3, 5, 7, 9, 11, 15]):
        f = time_batch(b2c.LINEAR_SAME, s, s, k, k, 65536, True, b2c.PATH_FFT, steps=3)
        d = time_batch(b2c.LINEAR_SAME, s, s, k, k, 65536, True, b2c.PATH_DIRECT, steps=3)
        emit("5: %dx%d (*) %dx%d LINEAR_SAME" % (s, s, k, k), {"fft_conv_per_s": f and f.get("conv_per_s"), "direct_conv_per_s": d and d.get("conv_per_s"),
                                                         "fft_period": f and f.get("fft"), "auto_path": b2c.Plan(b2c.LINEAR_SAME, s, s, k, k).info()["path"]})
