"""Run the trace variant (tools/build_variant.sh trace -DB2C_PFA_TRACE) on the headline workload and print the
per-pass clock profile of one image per CTA (median over all warps of 296 CTAs)."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["B2C_LIB_PATH"] = os.path.join(ROOT, "build", "variants", "trace.so")
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
G = int(os.environ.get("B2C_TRACE_CTAS", "296"))
N = G * 8 * 16
buf = (ctypes.c_longlong * N)()
L = b2c.lib()
L.b2c_debug_trace_read.argtypes = [ctypes.c_void_p, ctypes.c_long]
assert L.b2c_debug_trace_read(buf, N) == 0
a = np.array(buf, dtype=np.int64).reshape(G, 8, 16)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
np.save(os.path.join(ROOT, "gpurun_out", "trace_pfa108.npy"), a)
names = ["pass1", "bar", "pass2", "bar", "pass3", "bar", "pass4", "bar", "pass5"]
d = np.diff(a[:, :, :10], axis=2)
print("segment,median_cycles,mean_cycles,p90_cycles   (per warp; 2 CTAs share an SM)")
for i, nm in enumerate(names):
    print("%s,%.0f,%.0f,%.0f" % (nm, np.median(d[:, :, i]), d[:, :, i].mean(), np.percentile(d[:, :, i], 90)))
tot = a[:, :, 9] - a[:, :, 0]
print("image_total,%.0f,%.0f,%.0f" % (np.median(tot), tot.mean(), np.percentile(tot, 90)))
