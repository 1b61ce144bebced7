import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["B2C_LIB_PATH"] = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "build", "variants", "bigs_trace.so")
import torch, fftconvolution_b200 as b2c
torch.cuda.set_device(0)
n = 296 * 8
plan = b2c.Plan(b2c.CIRCULAR_FULL, 100, 100, 100, 100)
src = torch.rand((n, 100, 100), dtype=torch.float64, device="cuda"); ker = torch.rand_like(src)
dst = torch.empty((n, 199, 199), dtype=torch.float64, device="cuda")
plan.convolve_batch_device(src.data_ptr(), ker.data_ptr(), 0, dst.data_ptr(), n, 0); torch.cuda.synchronize()
buf = (ctypes.c_longlong * 16)()
L = b2c.lib(); L.b2c_debug_bigs_read.argtypes = [ctypes.c_void_p]; L.b2c_debug_bigs_read(buf)
names = ["R load", "R fft", "R unpack->slab", "C load", "C fft fwd", "C product", "C fft inv", "C store->slab", "R' load", "R' fft", "R' store", "C st fence+sync", "C st issue+wait.read", "C final wait", "R copies land", "R' copies land"]
tot = sum(buf[:16])
for i, nm in enumerate(names): print("%-16s %9.0f cycles/image  %5.1f%%" % (nm, buf[i] / 8, 100 * buf[i] / tot))
print("total %.0f cycles/image" % (tot / 8))
