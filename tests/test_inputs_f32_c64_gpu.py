"""single-precision and complex inputs (additive; the reference is double-real only,
convolution_fftw.h:216).  Both go through the C-ABI; the arithmetic is the FP64 real engine.
Tolerances: complex128 1e-12 relative L2 like every other parity test; float32 buffers 2e-7 (the result is rounded
to float once, 2^-24 per element; the inputs are exactly representable in double)."""
import numpy as np
import pytest

import _oracle

pytestmark = pytest.mark.gpu

CASES = [(2, 100, 100, 15, 15), (0, 30, 41, 7, 5), (3, 64, 64, 9, 9), (7, 20, 20, 20, 20), (4, 33, 35, 7, 9),
         (2, 1, 1000, 1, 63), (2, 12, 12, 3, 3), (5, 1, 1, 1, 1)]


@pytest.mark.parametrize("shared", [1, 0])
def test_float32_buffers(b2c, oracle, shared):
    import torch
    torch.cuda.set_device(0)
    rng = np.random.default_rng(71)
    for mode, hs, ws, hk, wk in CASES:
        for n in (1, 5):
            src = rng.random((n, hs, ws)).astype(np.float32)
            ker = (rng.random((hk, wk)) if shared else rng.random((n, hk, wk))).astype(np.float32)
            p = b2c.Plan(mode, hs, ws, hk, wk)
            d_src, d_ker = torch.from_numpy(src).cuda(), torch.from_numpy(ker).cuda()
            d_dst = torch.full((n, p.h_dst, p.w_dst), float("nan"), dtype=torch.float32, device="cuda")
            p.convolve_batch_device_f32(d_src.data_ptr(), d_ker.data_ptr(), shared, d_dst.data_ptr(), n, 0)
            torch.cuda.synchronize()
            got = d_dst.cpu().numpy()
            assert got.dtype == np.float32
            for i in (0, n - 1):
                want = oracle.convolve_direct8(mode, src[i].astype(np.float64), (ker if shared else ker[i]).astype(np.float64))
                assert _oracle.rel_err(got[i].astype(np.float64), want) <= 2e-7, (mode, hs, ws, hk, wk, n, shared, i)
            p.destroy()


@pytest.mark.parametrize("shared", [1, 0])
def test_complex128_buffers(b2c, oracle, shared):
    import torch
    torch.cuda.set_device(0)
    rng = np.random.default_rng(72)
    for mode, hs, ws, hk, wk in CASES:
        n = 3
        src = rng.random((n, hs, ws)) - 0.5 + 1j * (rng.random((n, hs, ws)) - 0.5)
        ker = (rng.random((hk, wk)) + 1j * rng.random((hk, wk))) if shared else (rng.random((n, hk, wk)) + 1j * rng.random((n, hk, wk)))
        p = b2c.Plan(mode, hs, ws, hk, wk)
        d_src, d_ker = torch.from_numpy(src).cuda(), torch.from_numpy(ker).cuda()
        d_dst = torch.full((n, p.h_dst, p.w_dst), complex("nan"), dtype=torch.complex128, device="cuda")
        p.convolve_batch_device_c64(d_src.data_ptr(), d_ker.data_ptr(), shared, d_dst.data_ptr(), n, 0)
        torch.cuda.synchronize()
        got = d_dst.cpu().numpy()
        for i in (0, n - 1):
            k = ker if shared else ker[i]
            c = lambda x, y: oracle.convolve_direct8(mode, np.ascontiguousarray(x), np.ascontiguousarray(y))
            want = (c(src[i].real, k.real) - c(src[i].imag, k.imag)) + 1j * (c(src[i].real, k.imag) + c(src[i].imag, k.real))
            err = np.linalg.norm((got[i] - want).ravel()) / np.linalg.norm(want.ravel())
            assert got[i].shape == want.shape and err <= 1e-12, (mode, hs, ws, hk, wk, shared, i, err)
        p.destroy()
