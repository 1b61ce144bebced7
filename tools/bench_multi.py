"""In-process multi-GPU host batch (b2c_multi_convolve_batch_host): headline shape from pinned host buffers."""
import sys, os, time, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, fftconvolution_b200 as b2c
n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
L = b2c.lib()
L.b2c_multi_convolve_batch_host.argtypes = [ctypes.c_int] * 5 + [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_long, ctypes.c_int]
src = torch.rand((n, 100, 100), dtype=torch.float64).pin_memory()
dst = torch.empty((n, 100, 100), dtype=torch.float64).pin_memory()
ker = torch.rand((15, 15), dtype=torch.float64).pin_memory()
ref = None
for nd in range(1, torch.cuda.device_count() + 1):
    for rep in range(2):
        t0 = time.perf_counter()
        rc = L.b2c_multi_convolve_batch_host(2, 100, 100, 15, 15, src.data_ptr(), ker.data_ptr(), 1, dst.data_ptr(), n, nd)
        dt = time.perf_counter() - t0
        assert rc == 0, b2c.lib().b2c_last_error()
    if ref is None: ref = dst.clone()
    print("devices %d: %.1f ms  %.0f k conv/s (incl. per-call plan + pipeline set-up)  identical to 1 device: %s" % (nd, dt * 1e3, n / dt / 1e3, bool(torch.equal(dst, ref))), flush=True)
