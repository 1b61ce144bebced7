"""Exploration: single-call latency through b2c_convolve_host (BASELINE configs[0]) with a pinned result
buffer, as the drop-in header owns one. Not a test."""
import sys, os, time, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, fftconvolution_b200 as b2c
L = b2c.lib()
L.b2c_convolve_host.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
rng = np.random.default_rng(1)
for (mode, hs, hk) in [(2, 50, 10), (2, 100, 15), (2, 30, 3), (2, 10, 3), (7, 100, 100)]:
    src, ker = rng.random((hs, hs)), rng.random((hk, hk))
    ker2 = rng.random((hk, hk))
    plan = b2c.Plan(mode, hs, hs, hk, hk)
    nb = plan.h_dst * plan.w_dst * 8
    hp = ctypes.c_void_p()
    assert L.b2c_host_alloc(ctypes.byref(hp), nb) == 0
    for forced in (b2c.PATH_AUTO, b2c.PATH_FFT, b2c.PATH_DIRECT):
        try:
            plan.set_path(forced)
            def call(k):
                rc = L.b2c_convolve_host(plan._h, src.ctypes.data, k.ctypes.data, hp)
                assert rc == 0, rc
            for _ in range(5): call(ker)
        except Exception as e:
            print("   path %d: %s" % (forced, e)); continue
        t0 = time.perf_counter()
        for _ in range(500): call(ker)
        same = (time.perf_counter() - t0) / 500
        t0 = time.perf_counter()
        for i in range(500): call(ker if i & 1 else ker2)
        alt = (time.perf_counter() - t0) / 500
        print("mode %d %dx%d (*) %dx%d forced path %d: %.1f us per call (same kernel), %.1f us (alternating kernels)" % (mode, hs, hs, hk, hk, forced, same * 1e6, alt * 1e6), flush=True)
    L.b2c_host_free(hp); plan.destroy()
