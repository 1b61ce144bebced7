"""1-D direct-kernel throughput: python tools/bench_direct1d.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, fftconvolution_b200 as b2c
torch.cuda.set_device(0)
tag = "nobulk" if os.environ.get("B2C_DIRECT_NO_BULK") else "bulk"
for mode, ns, nk in [(2, 4096, 3), (2, 4095, 3), (2, 4096, 8), (2, 4096, 15), (2, 4095, 15), (0, 1000, 5), (0, 999, 5), (2, 64, 15), (2, 63, 15), (2, 256, 3), (2, 255, 3), (3, 4096, 4)]:
    for col in (0, 1):
        shp = (ns, 1, nk, 1) if col else (1, ns, 1, nk)
        p = b2c.Plan(mode, *shp); p.set_path(b2c.PATH_DIRECT)
        n = int(max(2048, min(1 << 20, 4e8 / (8 * (ns + p.h_dst * p.w_dst)))))
        src = torch.rand((n, shp[0], shp[1]), dtype=torch.float64, device="cuda"); ker = torch.rand((shp[2], shp[3]), dtype=torch.float64, device="cuda")
        dst = torch.empty((n, p.h_dst, p.w_dst), dtype=torch.float64, device="cuda")
        for _ in range(3): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, 0)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, 0)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        gbs = n * 8 * (ns + p.h_dst * p.w_dst) / ms / 1e6
        print("%s: mode %d %dx%d (*) %dx%d: %.3f ms  %.2f M conv/s  %.0f GB/s = %.1f%% of roof" % (tag, mode, shp[0], shp[1], shp[2], shp[3], ms, n / ms / 1e3, gbs, gbs / 65.306))
        p.destroy()
