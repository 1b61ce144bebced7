"""Run the TRACE variant (build/variants/trace.so) and dump per-warp pass timelines of one image per CTA."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["B2C_LIB_PATH"] = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "build", "variants", "trace.so")
import numpy as np, torch
import fftconvolution_b200 as b2c
n = 65536
torch.cuda.set_device(0)
src = torch.empty((n, 100, 100), dtype=torch.float64, device="cuda"); dst = torch.empty_like(src)
b2c.fill_uniform_device(src.data_ptr(), src.numel(), 1)
ker = torch.from_numpy(np.random.default_rng(0).random((15, 15))).cuda()
plan = b2c.Plan(b2c.LINEAR_SAME, 100, 100, 15, 15)
for _ in range(2):
    plan.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, 0)
torch.cuda.synchronize()
N = 296 * 8 * 16
buf = (ctypes.c_longlong * N)()
L = b2c.lib()
L.b2c_debug_trace_read.argtypes = [ctypes.c_void_p, ctypes.c_long]
assert L.b2c_debug_trace_read(buf, N) == 0
a = np.array(buf, dtype=np.int64).reshape(296, 8, 16)
np.save("gpurun_out/trace_t108.npy", a)
# summary: per pass durations (median over CTAs/warps), barrier waits
names = ["p1", "bar", "p2", "bar", "p3", "bar", "p4", "bar", "p5", "bar", "p6", "bar", "p7"]
d = np.diff(a[:, :, :14], axis=2)
print("median cycles per segment (per warp):")
for i, nm in enumerate(names):
    print("  %-4s median %7.0f  mean %7.0f  p90 %7.0f" % (nm, np.median(d[:, :, i]), d[:, :, i].mean(), np.percentile(d[:, :, i], 90)))
tot = a[:, :, 13] - a[:, :, 0]
print("image total: median %.0f mean %.0f" % (np.median(tot), tot.mean()))
