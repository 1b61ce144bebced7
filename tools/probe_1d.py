"""Exploration: 1-D FFT path (tuned menu periods vs any-size kernel). Not a test."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, fftconvolution_b200 as b2c
torch.cuda.set_device(0)
st = torch.cuda.current_stream().cuda_stream
PEAK = 6540.2e9
rng = np.random.default_rng(3)
for (mode, ns, nk) in [(2, 63, 255), (2, 64, 64), (2, 100, 40), (0, 150, 30), (2, 200, 100), (2, 300, 33), (2, 400, 200), (2, 4096, 255)]:
    n = 16384
    src = torch.empty((n, 1, ns), dtype=torch.float64, device="cuda"); b2c.fill_uniform_device(src.data_ptr(), src.numel(), 5)
    ker = torch.rand((1, nk), dtype=torch.float64, device="cuda")
    out = []
    for path in (b2c.PATH_FFT, b2c.PATH_FFT_GENERIC):
        p = b2c.Plan(mode, 1, ns, 1, nk); p.set_path(path)
        dst = torch.empty((n, p.h_dst, p.w_dst), dtype=torch.float64, device="cuda")
        for _ in range(2): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, st)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5): p.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, st)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        out.append((ms, dst.clone(), p.w_fft, p.w_dst))
    b = 8 * (ns + out[0][3])
    err = (out[0][1] - out[1][1]).abs().max().item() / out[1][1].abs().max().item()
    print("mode %d 1x%d (*) 1x%d: P=%d tuned %.3f ms %.1f M conv/s %.1f%% of roof | any-size %.3f ms %.1f M conv/s | rel diff %.1e" % (
        mode, ns, nk, out[0][2], out[0][0], n / out[0][0] / 1e3, 100 * n / out[0][0] * 1e3 * b / PEAK, out[1][0], n / out[1][0] / 1e3, err), flush=True)
