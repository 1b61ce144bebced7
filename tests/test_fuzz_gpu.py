"""Seeded randomised parity sweep over BASELINE.json configs[4]'s grid (source 1..100 x kernel 1..100, all eight
modes, kernels larger than the source included) with the path left to the library (AUTO), batches of three,
shared and per-image kernels, and magnitudes spread over many decades -- against the oracle's long-double
nested loops (convolve_direct8).  Grown out of tools/fuzz_parity.py, which found two accuracy bugs in round 1
(DESIGN.md 5.1).  Categories: convolution_test.cc:424-523; tolerance 1e-12 relative L2."""
import os

import numpy as np
import pytest

import _oracle

pytestmark = pytest.mark.gpu

REL_TOL = 1e-12
POINTS = 320


def _points(seed):
    rng = np.random.default_rng(seed)
    out = []
    while len(out) < POINTS:
        kind = rng.random()
        if kind < 0.2:      # 1-D rows / columns
            ns, nk = int(rng.integers(1, 101)), int(rng.integers(1, 101))
            hs, ws, hk, wk = (1, ns, 1, nk) if rng.random() < 0.5 else (ns, 1, nk, 1)
        elif kind < 0.6:    # squares, the reference benchmark's grid (run_fftw_bench.sh:13-19) and beyond it
            ns, nk = int(rng.integers(1, 101)), int(rng.integers(1, 101))
            hs, ws, hk, wk = ns, ns, nk, nk
        else:               # rectangles
            hs, ws, hk, wk = (int(v) for v in rng.integers(1, 101, size=4))
        out.append((int(rng.integers(0, 8)), hs, ws, hk, wk, int(rng.random() < 0.7), int(rng.integers(0, 2 ** 31))))
    return out


@pytest.mark.parametrize("chunk", range(8))
def test_auto_path_matches_oracle_on_the_sweep_grid(b2c, oracle, chunk):
    import torch
    torch.cuda.set_device(0)
    pts = _points(int(os.environ.get("B2C_FUZZ_SEED", 20260)))[chunk * (POINTS // 8):(chunk + 1) * (POINTS // 8)]
    n = 3
    for mode, hs, ws, hk, wk, shared, seed in pts:
        rng = np.random.default_rng(seed)
        p = b2c.Plan(mode, hs, ws, hk, wk)
        hfft, wfft, hd, wd = oracle.plan(mode, hs, ws, hk, wk)
        assert (p.h_dst, p.w_dst) == (hd, wd), (mode, hs, ws, hk, wk)
        if p.h_dst == 0:      # impossible LINEAR_VALID: convolve is a no-op (convolution_fftw.h:218-219)
            p.destroy()
            continue
        s_scale, k_scale = 10.0 ** rng.uniform(-6, 6), 10.0 ** rng.uniform(-6, 6)
        src = s_scale * (rng.random((n, hs, ws)) - 0.3) * (10.0 ** rng.uniform(-4, 4, size=(n, 1, 1)))
        ker = k_scale * ((rng.random((hk, wk)) if shared else rng.random((n, hk, wk))) - 0.3)
        d_src, d_ker = torch.from_numpy(src).cuda(), torch.from_numpy(ker).cuda()
        d_dst = torch.full((n, hd, wd), float("nan"), dtype=torch.float64, device="cuda")
        p.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), shared, d_dst.data_ptr(), n, 0)
        torch.cuda.synchronize()
        got = d_dst.cpu().numpy()
        for i in (0, n - 1):
            want = oracle.convolve_direct8(mode, src[i], ker if shared else ker[i])
            err = _oracle.rel_err(got[i], want)
            assert got[i].shape == want.shape and err <= REL_TOL, (b2c.MODE_NAMES[mode], hs, ws, hk, wk, shared, i, err)
        p.destroy()


def _direct_points(seed, count):
    """Shapes for the forced direct path: kernels up to 16 rows (the register-window variants), images up to 240 rows
    (row tiles, double-buffered half images), even and odd widths (bulk-staged and element-wise staging), batches that
    give one unit, a partial wave, and several units per persistent CTA."""
    rng = np.random.default_rng(seed)
    out = []
    while len(out) < count:
        hk, wk = int(rng.integers(1, 17)), int(rng.integers(1, 17))
        big = rng.random() < 0.35
        hs = int(rng.integers(64, 241)) if big else int(rng.integers(1, 101))
        ws = int(rng.integers(64, 241)) if big else int(rng.integers(1, 101))
        if rng.random() < 0.6:
            ws += ws & 1            # even widths take the bulk-staged kernel
        mode = int(rng.integers(0, 8))
        n = int(rng.choice([1, 3, 37, 700])) if hs * ws <= 2500 else int(rng.choice([1, 3, 37, 300]))
        out.append((mode, hs, ws, hk, wk, n, int(rng.integers(0, 2 ** 31))))
    return out


@pytest.mark.parametrize("chunk", range(4))
def test_direct_path_matches_oracle_over_shapes_and_batches(b2c, oracle, chunk):
    import torch
    torch.cuda.set_device(0)
    # (B2C_FUZZ_SEED: one-off runs over other points)
    for mode, hs, ws, hk, wk, n, seed in _direct_points(int(os.environ.get("B2C_FUZZ_SEED", 77001)), 160)[chunk * 40:(chunk + 1) * 40]:
        rng = np.random.default_rng(seed)
        p = b2c.Plan(mode, hs, ws, hk, wk)
        if p.h_dst == 0 or p.w_dst == 0:
            p.destroy()
            continue
        p.set_path(b2c.PATH_DIRECT)
        hd, wd = p.h_dst, p.w_dst
        src = rng.random((n, hs, ws)) - 0.3
        ker = rng.random((hk, wk)) - 0.3
        d_src, d_ker = torch.from_numpy(src).cuda(), torch.from_numpy(ker).cuda()
        d_dst = torch.full((n, hd, wd), float("nan"), dtype=torch.float64, device="cuda")
        p.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), 1, d_dst.data_ptr(), n, 0)
        torch.cuda.synchronize()
        got = d_dst.cpu().numpy()
        assert not np.isnan(got).any(), ("unwritten outputs", b2c.MODE_NAMES[mode], hs, ws, hk, wk, n)
        for i in sorted({0, n // 2, n - 1}):
            want = oracle.convolve_direct8(mode, src[i], ker)
            err = _oracle.rel_err(got[i], want)
            assert err <= REL_TOL, (b2c.MODE_NAMES[mode], hs, ws, hk, wk, n, i, err)
        # every image of the batch against the first-principles identity: equal inputs give equal outputs
        if n >= 37:
            d_src[n - 2].copy_(d_src[1])
            p.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), 1, d_dst.data_ptr(), n, 0)
            torch.cuda.synchronize()
            assert torch.equal(d_dst[n - 2], d_dst[1]), ("unit position changes the result", hs, ws, hk, wk, n)
        p.destroy()
