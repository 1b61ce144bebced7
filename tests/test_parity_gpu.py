"""GPU parity tests: the CUDA path (through the C-ABI) against the CPU oracle on the same seeded
inputs, against the committed golden vectors, and -- at BASELINE.json's full sizes -- through
size-independent properties.  Tolerance: 1e-12 relative L2 (BASELINE.json north_star) and the
reference harness's 1e-10 absolute L2 (convolution_test.cc:49)."""
import os

import numpy as np
import pytest

import _oracle

pytestmark = pytest.mark.gpu

REL_TOL = 1e-12
ABS_TOL = 1e-10
GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "golden.npz"))
CASES = [tuple(int(v) for v in row) for row in GOLD["cases"]]


def _check(got, want, what):
    assert got.shape == want.shape, (what, got.shape, want.shape)
    if want.size == 0:
        return
    assert np.all(np.isfinite(got)), what
    rel = _oracle.rel_err(got, want)
    ab = float(np.linalg.norm((got - want).ravel()))
    assert rel <= REL_TOL, (what, rel)
    assert ab <= ABS_TOL * max(1.0, np.linalg.norm(want.ravel())), (what, ab)


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    torch.cuda.set_device(0)
    return torch


@pytest.mark.parametrize("ci", range(len(CASES)))
@pytest.mark.parametrize("path", ["fft", "direct"])
def test_single_call_matches_golden_all_modes(b2c, torch_cuda, ci, path):
    hs, ws, hk, wk, seed = CASES[ci]
    rng = np.random.default_rng(seed)
    src, ker = rng.random((hs, ws)), rng.random((hk, wk))
    for mode in range(8):
        key = "c%d_m%d" % (ci, mode)
        if key not in GOLD:
            continue
        p = b2c.Plan(mode, hs, ws, hk, wk)
        p.set_path(b2c.PATH_FFT if path == "fft" else b2c.PATH_DIRECT)
        try:
            got = p.convolve_host(src, ker)
        except b2c.B2CError as e:
            if path == "direct" and e.code == 4:   # kernel too large for the direct tile: documented limit
                p.destroy()
                continue
            raise
        _check(got, GOLD[key], (path, b2c.MODE_NAMES[mode], CASES[ci]))
        p.destroy()


def _random_case(rng, kind):
    # the reference test driver's nine size categories (convolution_test.cc:424-523)
    u = lambda: rng.random()
    if kind == 0: return 2 * int(20 * u() + 30), 1, 2 * int(20 * u() + 30), 1
    if kind == 1: return 2 * int(20 * u() + 30), 1, 2 * int(20 * u() + 20) + 1, 1
    if kind == 2: return 2 * int(20 * u() + 30) + 1, 1, int(20 * u() + 20), 1
    if kind == 3: return 2 * int(20 * u() + 30) + 1, 1, int(20 * u() + 20) + 1, 1
    if kind == 4: return 1, 2 * int(20 * u() + 30), 1, 2 * int(20 * u() + 20)
    if kind == 5: return 1, 2 * int(20 * u() + 30), 1, 2 * int(20 * u() + 20) + 1
    if kind == 6: return 1, 2 * int(20 * u() + 30) + 1, 1, int(20 * u() + 20)
    if kind == 7: return 1, 2 * int(20 * u() + 30) + 1, 1, int(20 * u() + 20) + 1
    return int(20 * u() + 30), int(20 * u() + 30), int(20 * u() + 20), int(20 * u() + 20)


@pytest.mark.parametrize("kind", range(9))
def test_reference_test_categories_vs_oracle(b2c, oracle, torch_cuda, kind):
    """Seeded re-run of the reference's 9 categories x all 8 modes (it only pins 6 of them against
    Octave); the oracle is the long-double nested-loop restatement."""
    rng = np.random.default_rng(100 + kind)
    for _ in range(6):
        hs, ws, hk, wk = _random_case(rng, kind)
        src, ker = rng.random((hs, ws)), rng.random((hk, wk))
        for mode in range(8):
            want = oracle.convolve_direct8(mode, src, ker)
            ws_ = b2c.Workspace()
            b2c.init_workspace(ws_, mode, hs, ws, hk, wk)
            b2c.convolve(ws_, src, ker)
            got = ws_.dst if ws_.dst is not None else np.zeros((0, 0))
            _check(got.reshape(want.shape) if want.size == 0 else got, want, (b2c.MODE_NAMES[mode], hs, ws, hk, wk))
            b2c.clear_workspace(ws_)


def test_edge_sizes_vs_oracle(b2c, oracle, torch_cuda):
    rng = np.random.default_rng(5)
    shapes = [(1, 1, 1, 1), (1, 2, 1, 1), (2, 1, 3, 1), (3, 3, 3, 3), (3, 5, 7, 2), (8, 8, 1, 1), (5, 4, 5, 4),
              (13, 11, 26, 23), (2, 100, 2, 99), (100, 2, 1, 2), (97, 89, 5, 3), (64, 64, 64, 64)]
    for hs, ws, hk, wk in shapes:
        src, ker = rng.random((hs, ws)), rng.random((hk, wk))
        for mode in range(8):
            want = oracle.convolve_direct8(mode, src, ker)
            for path in (b2c.PATH_FFT, b2c.PATH_DIRECT):
                p = b2c.Plan(mode, hs, ws, hk, wk)
                p.set_path(path)
                got = p.convolve_host(src, ker)
                _check(got, want, (b2c.MODE_NAMES[mode], path, hs, ws, hk, wk))
                p.destroy()


def test_batch_device_shared_kernel_equals_single_calls(b2c, oracle, torch_cuda):
    torch = torch_cuda
    rng = np.random.default_rng(11)
    for (hs, ws, hk, wk, mode, n) in [(50, 50, 10, 10, 2, 7), (100, 100, 15, 15, 2, 10), (33, 20, 12, 9, 0, 5),
                                      (41, 1, 8, 1, 4, 9), (1, 77, 1, 30, 3, 4)]:
        src = rng.random((n, hs, ws))
        ker = rng.random((hk, wk))
        for path in (b2c.PATH_FFT, b2c.PATH_DIRECT):
            p = b2c.Plan(mode, hs, ws, hk, wk)
            p.set_path(path)
            d_src = torch.from_numpy(src).cuda()
            d_ker = torch.from_numpy(ker).cuda()
            d_dst = torch.full((n, p.h_dst, p.w_dst), float("nan"), dtype=torch.float64, device="cuda")
            p.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), 1, d_dst.data_ptr(), n,
                                    torch.cuda.current_stream().cuda_stream)
            torch.cuda.synchronize()
            got = d_dst.cpu().numpy()
            for i in range(n):
                _check(got[i], oracle.convolve_direct8(mode, src[i], ker), (path, mode, i))
            p.destroy()


def test_batch_device_per_image_kernels(b2c, oracle, torch_cuda):
    torch = torch_cuda
    rng = np.random.default_rng(12)
    for (hs, ws, hk, wk, mode, n) in [(20, 20, 20, 20, 7, 5), (30, 25, 7, 9, 2, 4), (17, 1, 40, 1, 4, 3),
                                      (1, 50, 1, 50, 0, 3), (25, 31, 6, 4, 3, 3), (100, 100, 100, 100, 7, 2)]:
        src = rng.random((n, hs, ws))
        ker = rng.random((n, hk, wk))
        p = b2c.Plan(mode, hs, ws, hk, wk)
        p.set_path(b2c.PATH_FFT)
        d_src = torch.from_numpy(src).cuda()
        d_ker = torch.from_numpy(ker).cuda()
        d_dst = torch.full((n, p.h_dst, p.w_dst), float("nan"), dtype=torch.float64, device="cuda")
        p.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), 0, d_dst.data_ptr(), n, 0)
        torch.cuda.synchronize()
        got = d_dst.cpu().numpy()
        for i in range(n):
            _check(got[i], oracle.convolve_direct8(mode, src[i], ker[i]), (mode, i, hs, ws, hk, wk))
        p.destroy()


def test_config3_shape_and_circular_full_identity(b2c, oracle, torch_cuda):
    """BASELINE config 3: 100x100 (*) 100x100 CIRCULAR_FULL, per-image kernels -> 199x199, equal to
    LINEAR_FULL (period == full length)."""
    rng = np.random.default_rng(13)
    src, ker = rng.random((100, 100)), rng.random((100, 100))
    a = b2c.Plan(b2c.CIRCULAR_FULL, 100, 100, 100, 100)
    assert (a.h_dst, a.w_dst) == (199, 199)
    got = a.convolve_host(src, ker)
    _check(got, oracle.convolve_fftw(b2c.CIRCULAR_FULL, src, ker), "config3 vs reference FFT path restated")
    b = b2c.Plan(b2c.LINEAR_FULL, 100, 100, 100, 100)
    _check(b.convolve_host(src, ker), got, "LINEAR_FULL == CIRCULAR_FULL")


def test_strip_mode_many_images_per_cta(b2c, oracle, torch_cuda):
    """Strip-mode kernels are persistent: one CTA walks several images through the same slab and the
    same shared-memory buffers.  A batch much larger than the grid, built from a few distinct
    (source, kernel) pairs, must give every copy the bit pattern of its prototype (which is checked
    against the oracle) -- catches stale slab / cache lines carried from one image to the next."""
    torch = torch_cuda
    rng = np.random.default_rng(77)
    # the last two also run the compile-time 200 x 200 kernel: odd widths (rows loaded element-wise) and
    # source / kernel of different heights with even widths (rows staged by bulk copies)
    for (mode, hs, ws, hk, wk, n) in [(7, 100, 100, 100, 100, 2 * 148 + 41), (0, 120, 90, 70, 100, 148 + 19),
                                      (7, 101, 99, 100, 101, 2 * 148 + 30), (0, 130, 150, 71, 50, 2 * 148 + 10),
                                      (0, 150, 100, 80, 60, 148 + 23),    # 229 x 159: runtime-schedule strip kernel
                                      # CIRCULAR_SAME of non-smooth sizes by periodic extension on the 200 x 200 kernel,
                                      # kernel larger than / smaller than the source
                                      (4, 97, 97, 99, 99, 148 + 20), (4, 89, 93, 80, 85, 148 + 5)]:
        nproto = 3
        psrc, pker = rng.random((nproto, hs, ws)), rng.random((nproto, hk, wk))
        which = rng.integers(0, nproto, size=n)
        p = b2c.Plan(mode, hs, ws, hk, wk)
        assert mode == 4 or not p.info()["work_in_smem"]
        p.set_path(b2c.PATH_FFT)
        d_src = torch.from_numpy(psrc[which]).cuda()
        d_ker = torch.from_numpy(pker[which]).cuda()
        d_dst = torch.full((n, p.h_dst, p.w_dst), float("nan"), dtype=torch.float64, device="cuda")
        for _ in range(2):   # second pass starts from slabs left by the first
            p.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), 0, d_dst.data_ptr(), n, 0)
        torch.cuda.synchronize()
        got = d_dst.cpu().numpy()
        first = [int(np.flatnonzero(which == j)[0]) for j in range(nproto)]
        for j in range(nproto):
            _check(got[first[j]], oracle.convolve_direct8(mode, psrc[j], pker[j]), ("prototype", mode, j))
        for i in range(n):
            assert np.array_equal(got[i], got[first[which[i]]]), ("copy differs from its prototype", mode, i)
        p.destroy()


def test_strip_mode_split_form_small_batches_and_large_images(b2c, oracle, torch_cuda):
    """Batches far from a whole number of waves run one launch per phase with one CTA per chunk / strip
    (conv_big2d.cu split form): every crop mode, rectangles, shared and per-image kernels, batch sizes
    across the group boundary (74 on 148 SMs), one 1024 x 1024 and one 4096 x 4096 image."""
    torch = torch_cuda
    rng = np.random.default_rng(91)
    cases = [(0, 150, 130, 31, 60, 1, 1), (2, 170, 90, 40, 120, 5, 0), (3, 160, 170, 9, 30, 2, 1), (5, 128, 128, 64, 33, 3, 0),
             (7, 100, 100, 100, 100, 75, 0), (2, 200, 190, 17, 17, 160, 1), (2, 1024, 1024, 33, 33, 1, 1),
             (2, 4096, 4096, 5, 5, 1, 1)]   # last: beyond the any-size kernel's index range, strip mode only
    for (mode, hs, ws, hk, wk, n, shared) in cases:
        p = b2c.Plan(mode, hs, ws, hk, wk)
        assert not p.info()["work_in_smem"], (mode, hs, ws, hk, wk)
        p.set_path(b2c.PATH_FFT)
        src = rng.random((n, hs, ws))
        ker = rng.random((hk, wk)) if shared else rng.random((n, hk, wk))
        d_src, d_ker = torch.from_numpy(src).cuda(), torch.from_numpy(ker).cuda()
        d_dst = torch.full((n, p.h_dst, p.w_dst), float("nan"), dtype=torch.float64, device="cuda")
        p.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), shared, d_dst.data_ptr(), n, 0)
        torch.cuda.synchronize()
        got = d_dst.cpu().numpy()
        for i in sorted({0, n // 2, n - 1}):
            want = oracle.convolve_direct8(mode, src[i], ker if shared else ker[i])
            _check(got[i], want, ("split form", mode, hs, ws, hk, wk, n, i))
        assert np.isfinite(got).all()
        p.destroy()


def test_tuned_1d_period_family_vs_oracle(b2c, oracle, torch_cuda):
    """1-D signals with a shared kernel on every menu period of conv_tuned1d.cu (three- and four-pass
    schedules), rows and columns, all crop modes, odd batches (the last CTA transforms one signal), against
    the oracle and against the any-size kernel."""
    torch = torch_cuda
    rng = np.random.default_rng(53)
    cases = [(2, 100, 40, 128), (0, 150, 30, 192), (2, 63, 255, 192), (3, 256, 100, 256), (2, 300, 33, 320), (0, 301, 70, 384),
             (2, 480, 64, 512), (0, 600, 41, 640), (3, 768, 100, 768), (2, 1000, 40, 1024), (0, 1000, 255, 1280), (5, 1200, 300, 1536),
             (2, 2000, 64, 2048), (0, 2100, 200, 2304), (2, 2500, 120, 2560), (7, 3000, 73, 3072), (4, 4096, 255, 4096),
             (2, 4096, 255, 4608), (0, 4096, 1000, 5120), (2, 4096, 4096, 6144), (0, 4095, 4095, 8192), (6, 4000, 97, 4096)]
    for k, (mode, ns, nk, P) in enumerate(cases):
        column = k % 3 == 2
        shape_s, shape_k = ((ns, 1), (nk, 1)) if column else ((1, ns), (1, nk))
        p = b2c.Plan(mode, *shape_s, *shape_k)
        assert (p.h_fft if column else p.w_fft) == P, (mode, ns, nk, p.h_fft, p.w_fft, P)
        n = 5
        src, ker = rng.random((n,) + shape_s), rng.random(shape_k)
        d_src, d_ker = torch.from_numpy(src).cuda(), torch.from_numpy(ker).cuda()
        got = {}
        for path in (b2c.PATH_FFT, b2c.PATH_FFT_GENERIC):
            p.set_path(path)
            d_dst = torch.full((n, p.h_dst, p.w_dst), float("nan"), dtype=torch.float64, device="cuda")
            p.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), 1, d_dst.data_ptr(), n, 0)
            torch.cuda.synchronize()
            got[path] = d_dst.cpu().numpy()
        for i in (0, n - 1):
            _check(got[b2c.PATH_FFT][i], oracle.convolve_direct8(mode, src[i], ker), ("tuned 1-D", mode, ns, nk, P, i))
        _check(got[b2c.PATH_FFT], got[b2c.PATH_FFT_GENERIC], ("tuned 1-D vs any-size", mode, ns, nk, P))
        p.destroy()


def test_single_call_kernel_cache(b2c, oracle, torch_cuda):
    """convolve() through the header re-uses the uploaded kernel and its spectrum while the caller passes the
    same kernel contents (the reference benchmark's loop), and must notice every change: new contents in the
    same buffer, a batch call with another kernel in between, a forced path change."""
    torch = torch_cuda
    rng = np.random.default_rng(61)
    for (mode, hs, ws, hk, wk) in [(2, 50, 50, 10, 10), (2, 100, 100, 15, 15), (0, 1, 4096, 1, 255), (2, 30, 30, 3, 3), (7, 100, 100, 100, 100),
                                   (4, 33, 35, 7, 9)]:
        p = b2c.Plan(mode, hs, ws, hk, wk)
        a, b = rng.random((hs, ws)), rng.random((hs, ws))
        k1, k2 = rng.random((hk, wk)), rng.random((hk, wk))
        want = lambda s, k: oracle.convolve_direct8(mode, s, k)
        _check(p.convolve_host(a, k1), want(a, k1), ("first call", mode))
        _check(p.convolve_host(b, k1), want(b, k1), ("cached kernel, new source", mode))
        kbuf = k1.copy()
        _check(p.convolve_host(a, kbuf), want(a, k1), ("equal contents in another buffer", mode))
        kbuf[...] = k2
        _check(p.convolve_host(a, kbuf), want(a, k2), ("same buffer, new contents", mode))
        # a batch call with a third kernel overwrites the plan's spectrum
        k3 = rng.random((hk, wk))
        d_src, d_ker = torch.from_numpy(np.stack([a, b])).cuda(), torch.from_numpy(k3).cuda()
        d_dst = torch.empty((2, p.h_dst, p.w_dst), dtype=torch.float64, device="cuda")
        p.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), 1, d_dst.data_ptr(), 2, 0)
        torch.cuda.synchronize()
        _check(d_dst[1].cpu().numpy(), want(b, k3), ("batch call", mode))
        _check(p.convolve_host(b, kbuf), want(b, k2), ("single call after a batch call with another kernel", mode))
        for path in (b2c.PATH_FFT_GENERIC, b2c.PATH_FFT):
            p.set_path(path)
            _check(p.convolve_host(a, kbuf), want(a, k2), ("after set_path", mode, path))
        p.destroy()


def test_multi_device_batch_in_one_process(b2c, oracle, torch_cuda):
    """b2c_multi_convolve_batch_host: contiguous shards of a host batch, one host thread + plan + pipeline per
    device of THIS process.  Runs with however many GPUs are visible (1 on the test box: the shard logic is
    also covered on CPU by tests/test_partition_gloo.py); must equal the single-plan batch (bit for bit on one device)."""
    torch = torch_cuda
    rng = np.random.default_rng(71)
    ndev = torch.cuda.device_count()
    for (mode, hs, ws, hk, wk, n, shared) in [(2, 40, 36, 7, 9, 37, 1), (7, 20, 20, 20, 20, 11, 0), (0, 1, 700, 1, 90, 9, 1)]:
        src = rng.random((n, hs, ws))
        ker = rng.random((hk, wk)) if shared else rng.random((n, hk, wk))
        p = b2c.Plan(mode, hs, ws, hk, wk)
        want = p.convolve_batch_host(src, ker)
        p.destroy()
        for nd in sorted({0, 1, ndev}):
            got = b2c.multi_convolve_batch_host(mode, src, ker, n_devices=nd)
            if (ndev if nd == 0 else nd) == 1:
                assert np.array_equal(got, want), (mode, nd)
            else:   # a shard that starts on an odd image pairs images differently in the two-per-transform kernels
                _check(got, want, ("multi-device vs one plan", mode, nd))
        _check(want[n - 1], oracle.convolve_direct8(mode, src[n - 1], ker if shared else ker[n - 1]), ("multi", mode))
    with pytest.raises(b2c.B2CError) as e:
        b2c.multi_convolve_batch_host(2, np.zeros((2, 8, 8)), np.zeros((3, 3)), n_devices=ndev + 1)
    assert e.value.code == 1 and "devices requested" in str(e.value)
    torch.cuda.set_device(0)


def test_circular_same_on_tuned_menu_periods(b2c, oracle, torch_cuda):
    """CIRCULAR_SAME sizes whose own period has no tuned kernel (not smooth, or smooth but off the menu) run
    as the linear convolution of the periodically extended source at a menu period (planner.h
    make_plan_circular_ext): prime and odd sizes, rectangles, kernel == source, 1-D rows and columns, odd
    batches -- against the oracle and against the any-size kernel (exact period or fold path)."""
    torch = torch_cuda
    rng = np.random.default_rng(83)
    cases = [(31, 37, 5, 7), (97, 89, 9, 3), (45, 45, 45, 45), (63, 21, 1, 21), (100, 100, 9, 9), (50, 75, 11, 1),
             (1, 1000, 1, 255), (1, 997, 1, 64), (613, 1, 100, 1), (1, 63, 1, 63), (1, 3000, 1, 17)]
    for (hs, ws, hk, wk) in cases:
        p = b2c.Plan(b2c.CIRCULAR_SAME, hs, ws, hk, wk)
        assert (p.h_dst, p.w_dst) == (hs, ws)
        n = 5
        src, ker = rng.random((n, hs, ws)), rng.random((hk, wk))
        d_src, d_ker = torch.from_numpy(src).cuda(), torch.from_numpy(ker).cuda()
        got = {}
        for path in (b2c.PATH_FFT, b2c.PATH_FFT_GENERIC):
            p.set_path(path)
            d_dst = torch.full((n, hs, ws), float("nan"), dtype=torch.float64, device="cuda")
            p.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), 1, d_dst.data_ptr(), n, 0)
            torch.cuda.synchronize()
            got[path] = d_dst.cpu().numpy()
        for i in (0, n - 1):
            _check(got[b2c.PATH_FFT][i], oracle.convolve_direct8(b2c.CIRCULAR_SAME, src[i], ker), ("circular ext", hs, ws, hk, wk, i))
        _check(got[b2c.PATH_FFT], got[b2c.PATH_FFT_GENERIC], ("circular ext vs any-size", hs, ws, hk, wk))
        p.set_path(b2c.PATH_AUTO)
        _check(p.convolve_host(src[1], ker), got[b2c.PATH_FFT][1], ("single call", hs, ws, hk, wk))
        p.destroy()


def test_packed_transform_kernels_keep_accuracy_when_magnitudes_differ(b2c, oracle, torch_cuda):
    """Kernels that transform source and kernel as ONE complex image (per-image kernels; strip mode) extract the
    kernel spectrum as a difference of source-sized numbers: with images ~1e3 and kernels ~1e-4 the relative error
    was ~1e-9 until the kernel part got a power-of-two scale that equalises the magnitudes (balance_kernel).  The
    reference transforms the two separately (convolution_fftw.h:190-191); the 1e-12 bound must hold for any scaling."""
    torch = torch_cuda
    rng = np.random.default_rng(97)
    cases = [(2, 40, 44, 9, 7, 5, 0),          # per-image kernels, work buffer in shared memory
             (7, 100, 100, 100, 100, 3, 0),    # period 200: split form of the strip kernel (3 images)
             (7, 100, 100, 100, 100, 300, 0),  # ... persistent form (conv_big2ds.cu)
             (0, 150, 130, 31, 60, 2, 1),      # strip mode, shared kernel, split form
             (2, 170, 90, 40, 120, 160, 1)]    # strip mode, shared kernel, persistent form
    for (mode, hs, ws, hk, wk, n, shared) in cases:
        for (s_scale, k_scale) in [(1e3, 1e-4), (1e-5, 3e2)]:
            src = s_scale * (0.5 + rng.random((n, hs, ws)))
            ker = k_scale * (rng.random((hk, wk)) if shared else rng.random((n, hk, wk)))
            if not shared:
                ker[n - 1] *= 1e-3            # scales differ between images of one batch
            p = b2c.Plan(mode, hs, ws, hk, wk)
            p.set_path(b2c.PATH_FFT)
            d_src, d_ker = torch.from_numpy(src).cuda(), torch.from_numpy(ker).cuda()
            d_dst = torch.full((n, p.h_dst, p.w_dst), float("nan"), dtype=torch.float64, device="cuda")
            p.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), shared, d_dst.data_ptr(), n, 0)
            torch.cuda.synchronize()
            got = d_dst.cpu().numpy()
            for i in sorted({0, n - 1}):
                _check(got[i], oracle.convolve_direct8(mode, src[i], ker if shared else ker[i]),
                       ("magnitudes", mode, hs, ws, hk, wk, n, shared, s_scale, k_scale, i))
            p.destroy()
    # two signals per transform (1-D family): neighbours of a batch that differ by 1e9 in magnitude
    for (ns, nk) in [(1000, 255), (4096, 63), (100, 40), (4096, 4096), (1000, -255), (77, -30)]:   # negative: any-size kernel
        generic = nk < 0
        nk = abs(nk)
        n = 6
        src = rng.random((n, 1, ns)) - 0.5
        src[1] *= 1e9; src[2] *= 1e-7; src[5] *= 1e5
        ker = rng.random((1, nk))
        p = b2c.Plan(b2c.LINEAR_SAME, 1, ns, 1, nk)
        p.set_path(b2c.PATH_FFT_GENERIC if generic else b2c.PATH_FFT)
        d_src, d_ker = torch.from_numpy(src).cuda(), torch.from_numpy(ker).cuda()
        d_dst = torch.full((n, p.h_dst, p.w_dst), float("nan"), dtype=torch.float64, device="cuda")
        p.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), 1, d_dst.data_ptr(), n, 0)
        torch.cuda.synchronize()
        got = d_dst.cpu().numpy()
        for i in range(n):
            _check(got[i], oracle.convolve_direct8(b2c.LINEAR_SAME, src[i], ker), ("pair magnitudes", ns, nk, i, generic))
        p.destroy()


def test_one_d_up_to_4096(b2c, oracle, torch_cuda):
    """BASELINE config 4 at sizes the oracle finishes in seconds: odd/even source x odd/even kernel."""
    rng = np.random.default_rng(14)
    for n, m in [(63, 16), (64, 15), (255, 256), (1000, 3), (4095, 4), (4096, 255), (4096, 4097)]:
        src, ker = rng.random((1, n)), rng.random((1, m))
        for mode in range(8):
            want = oracle.convolve_direct8(mode, src, ker)
            p = b2c.Plan(mode, 1, n, 1, m)
            _check(p.convolve_host(src, ker), want, ("1xN", mode, n, m))
            p.destroy()
            p = b2c.Plan(mode, n, 1, m, 1)
            _check(p.convolve_host(src.T.copy(), ker.T.copy()), want.T, ("Nx1", mode, n, m))
            p.destroy()


def test_one_d_per_image_kernels_and_beyond_the_menu(b2c, oracle, torch_cuda):
    """1-D rows and columns with PER-SIGNAL kernels (batch), and a 1 x 20000 (*) 1 x 20000 LINEAR_FULL plan whose period is
    beyond both the tuned 1-D menu and strip mode: it runs on the any-size kernel with its work buffer in global memory."""
    torch = torch_cuda
    rng = np.random.default_rng(16)
    n = 5
    for mode, ns, nk in [(0, 257, 30), (2, 1000, 255), (3, 4096, 100), (4, 777, 64), (7, 4095, 256), (5, 64, 63)]:
        for column in (False, True):
            shape_s, shape_k = ((ns, 1), (nk, 1)) if column else ((1, ns), (1, nk))
            src, ker = rng.random((n,) + shape_s), rng.random((n,) + shape_k)
            p = b2c.Plan(mode, *shape_s, *shape_k)
            d_src, d_ker = torch.from_numpy(src).cuda(), torch.from_numpy(ker).cuda()
            d_dst = torch.full((n, p.h_dst, p.w_dst), float("nan"), dtype=torch.float64, device="cuda")
            p.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), 0, d_dst.data_ptr(), n, 0)
            torch.cuda.synchronize()
            got = d_dst.cpu().numpy()
            for i in (0, n - 1):
                _check(got[i], oracle.convolve_direct8(mode, src[i], ker[i]), ("1-D per-signal kernels", mode, ns, nk, column, i))
            p.destroy()
    src, ker = rng.random((1, 20000)), rng.random((1, 20000))
    p = b2c.Plan(b2c.LINEAR_FULL, 1, 20000, 1, 20000)
    _check(p.convolve_host(src, ker), oracle.convolve_direct8(b2c.LINEAR_FULL, src, ker), "1 x 20000 (*) 1 x 20000")
    p.destroy()


def test_host_batch_pipeline_matches_device_batch(b2c, oracle, torch_cuda):
    rng = np.random.default_rng(15)
    n, hs, ws, hk, wk = 1500, 100, 100, 15, 15   # > one pipeline chunk
    src = rng.random((n, hs, ws))
    ker = rng.random((hk, wk))
    p = b2c.Plan(b2c.LINEAR_SAME, hs, ws, hk, wk)
    got = p.convolve_batch_host(src, ker)
    for i in [0, 1, 2, 599, 600, 1198, 1199, 1499]:
        _check(got[i], oracle.convolve_direct8(b2c.LINEAR_SAME, src[i], ker), ("host pipeline", i))
    # linearity over the whole batch: conv(sum_i a_i) == sum_i conv(a_i)
    total = p.convolve_host(src.sum(axis=0), ker)
    assert _oracle.rel_err(got.sum(axis=0), total) <= 1e-11
    pk = rng.random((n, hk, wk))
    got2 = p.convolve_batch_host(src[:40], pk[:40])
    for i in [0, 17, 39]:
        _check(got2[i], oracle.convolve_direct8(b2c.LINEAR_SAME, src[i], pk[i]), ("host per-image", i))


def test_headline_full_size_properties(b2c, oracle, torch_cuda):
    """BASELINE config 2 at full size (65536 x 100x100 (*) 15x15 LINEAR_SAME, device resident):
    spot-check images against the oracle, commutation with a delta kernel, linearity, and equality of
    the checksum of the batch with the checksum of its two halves run separately."""
    torch = torch_cuda
    n, hs, ws, hk, wk = 65536, 100, 100, 15, 15
    seed = 0x5EED0002
    d_src = torch.empty((n, hs, ws), dtype=torch.float64, device="cuda")
    b2c.fill_uniform_device(d_src.data_ptr(), d_src.numel(), seed)
    torch.cuda.synchronize()
    host_first = d_src[:3].cpu().numpy()
    assert np.array_equal(host_first.ravel(), _oracle.splitmix_uniform(seed, 0, 3 * hs * ws))
    ker = np.random.default_rng(16).random((hk, wk))
    d_ker = torch.from_numpy(ker).cuda()
    d_dst = torch.empty((n, hs, ws), dtype=torch.float64, device="cuda")
    p = b2c.Plan(b2c.LINEAR_SAME, hs, ws, hk, wk)
    p.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), 1, d_dst.data_ptr(), n, 0)
    torch.cuda.synchronize()
    for i in [0, 1, 2, 12345, 65534, 65535]:
        img = _oracle.splitmix_uniform(seed, i * hs * ws, hs * ws).reshape(hs, ws)
        _check(d_dst[i].cpu().numpy(), oracle.convolve_direct8(b2c.LINEAR_SAME, img, ker), ("headline", i))
    whole = b2c.checksum_device(d_dst.data_ptr(), d_dst.numel())
    # halves, as two ranks would own them: checksums add up
    b0, e0 = b2c.partition(n, 2, 0)
    b1, e1 = b2c.partition(n, 2, 1)
    d_half = torch.empty((e0 - b0, hs, ws), dtype=torch.float64, device="cuda")
    parts = []
    for b, e in [(b0, e0), (b1, e1)]:
        p.convolve_batch_device(d_src[b:e].data_ptr(), d_ker.data_ptr(), 1, d_half.data_ptr(), e - b, 0)
        torch.cuda.synchronize()
        parts.append(b2c.checksum_device(d_half.data_ptr(), (e - b) * hs * ws))
    assert (parts[0][0] + parts[1][0]) % (1 << 64) == whole[0]
    assert parts[0][1] ^ parts[1][1] == whole[1]
    # delta kernel centred like LINEAR_SAME's crop: identity
    delta = np.zeros((hk, wk)); delta[hk // 2, wk // 2] = 1.0
    d_delta = torch.from_numpy(delta).cuda()
    p.convolve_batch_device(d_src.data_ptr(), d_delta.data_ptr(), 1, d_dst.data_ptr(), n, 0)
    torch.cuda.synchronize()
    assert float((d_dst - d_src).abs().max()) < 1e-13


@pytest.mark.parametrize("kernel", ["t2", "pfa"])
def test_tuned_108_kernel_vs_oracle_and_generic(b2c, oracle, torch_cuda, kernel, monkeypatch):
    """Plans whose periods are 108 x 108 (the headline shape) run on the (12, 9) member of conv_tuned2.cu; the
    five-pass prime-factor kernel of conv_pfa108.cu is the alternative behind B2C_P108=pfa.  Both against the
    oracle, and the batch bit-compared against itself run image by image."""
    torch = torch_cuda
    monkeypatch.setenv("B2C_P108", kernel)
    rng = np.random.default_rng(21)
    cases = [(2, 100, 100, 15, 15), (0, 90, 95, 19, 14), (3, 108, 108, 5, 5), (2, 101, 104, 14, 9),
             (0, 50, 60, 59, 49), (5, 100, 100, 9, 9), (7, 54, 54, 55, 55), (1, 100, 100, 16, 17),
             (2, 30, 100, 157, 17), (4, 97, 89, 9, 3)]
    for mode, hs, ws, hk, wk in cases:
        p = b2c.Plan(mode, hs, ws, hk, wk)
        if mode != b2c.CIRCULAR_SAME:   # (CIRCULAR_SAME reports its own period; it runs as a periodic extension at 108)
            assert (p.h_fft, p.w_fft) == (108, 108), (mode, hs, ws, hk, wk, p.h_fft, p.w_fft)
        p.set_path(b2c.PATH_FFT)
        n = 5
        src = rng.random((n, hs, ws))
        ker = rng.random((hk, wk))
        d_src, d_ker = torch.from_numpy(src).cuda(), torch.from_numpy(ker).cuda()
        d_dst = torch.full((n, p.h_dst, p.w_dst), float("nan"), dtype=torch.float64, device="cuda")
        p.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), 1, d_dst.data_ptr(), n, 0)
        torch.cuda.synchronize()
        got = d_dst.cpu().numpy()
        for i in range(n):
            _check(got[i], oracle.convolve_direct8(mode, src[i], ker), ("period 108", mode, hs, ws, hk, wk, i))
        g = b2c.Plan(mode, hs, ws, hk, wk)
        g.set_path(b2c.PATH_FFT_GENERIC)
        _check(g.convolve_host(src[0], ker), got[0], ("generic == tuned", mode))
        one = p.convolve_host(src[3], ker)
        assert np.array_equal(one, got[3]), "batch result must not depend on batch position"


def test_batch_edges_and_errors(b2c, oracle, torch_cuda):
    """Empty batch, batch of one, odd batch tails on every kernel family, and argument errors."""
    torch = torch_cuda
    rng = np.random.default_rng(31)
    L = b2c.lib()
    for (mode, hs, ws, hk, wk, path) in [(2, 100, 100, 15, 15, b2c.PATH_FFT), (0, 20, 30, 5, 7, b2c.PATH_FFT_GENERIC),
                                         (2, 40, 40, 5, 5, b2c.PATH_DIRECT), (7, 120, 90, 110, 130, b2c.PATH_FFT)]:
        p = b2c.Plan(mode, hs, ws, hk, wk)
        p.set_path(path)
        for n in (0, 1, 3):
            src = rng.random((max(n, 1), hs, ws)); ker = rng.random((hk, wk))
            d_src, d_ker = torch.from_numpy(src).cuda(), torch.from_numpy(ker).cuda()
            d_dst = torch.full((max(n, 1), p.h_dst, p.w_dst), -7.0, dtype=torch.float64, device="cuda")
            p.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), 1, d_dst.data_ptr(), n, 0)
            torch.cuda.synchronize()
            got = d_dst.cpu().numpy()
            for i in range(n):
                _check(got[i], oracle.convolve_direct8(mode, src[i], ker), ("edge batch", mode, n, i))
            assert np.all(got[n:] == -7.0), "images beyond the batch must stay untouched"
        import ctypes
        assert L.b2c_convolve_batch_device(p._h, None, None, 1, None, 4, None) == 1      # null buffers
        assert L.b2c_convolve_batch_device(p._h, None, None, 1, None, -1, None) == 1     # negative batch
        p.destroy()
    assert L.b2c_convolve_host(None, None, None, None) == 1
    assert b"null plan" in L.b2c_last_error()


def test_strip_mode_shared_kernel_and_rectangles(b2c, oracle, torch_cuda):
    """Transforms larger than shared memory (conv_big2d.cu) with a shared kernel, rectangular sizes
    and every crop mode; CIRCULAR_SAME with a non-smooth size falls back to the slab path (fold)."""
    torch = torch_cuda
    rng = np.random.default_rng(32)
    for (mode, hs, ws, hk, wk) in [(0, 150, 130, 31, 60), (2, 170, 90, 40, 120), (3, 160, 170, 9, 30), (5, 128, 128, 64, 33),
                                   (4, 131, 127, 20, 25)]:
        p = b2c.Plan(mode, hs, ws, hk, wk)
        assert not p.info()["work_in_smem"], (mode, hs, ws, hk, wk, p.h_fft, p.w_fft)
        p.set_path(b2c.PATH_FFT)
        n = 3
        src, ker = rng.random((n, hs, ws)), rng.random((hk, wk))
        d_src, d_ker = torch.from_numpy(src).cuda(), torch.from_numpy(ker).cuda()
        d_dst = torch.full((n, p.h_dst, p.w_dst), float("nan"), dtype=torch.float64, device="cuda")
        p.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), 1, d_dst.data_ptr(), n, 0)
        torch.cuda.synchronize()
        got = d_dst.cpu().numpy()
        for i in range(n):
            _check(got[i], oracle.convolve_direct8(mode, src[i], ker), ("strip mode", mode, hs, ws, hk, wk, i))
        p.destroy()


def test_tuned_square_period_family_vs_oracle(b2c, oracle, torch_cuda):
    """conv_tuned2.cu: one fused kernel per square period of the menu (16 ... 160).  Each period is hit
    with several modes / rectangular shapes and checked against the oracle and the any-size kernel."""
    torch = torch_cuda
    rng = np.random.default_rng(41)
    menu = [16, 20, 24, 28, 32, 36, 40, 42, 48, 50, 54, 56, 60, 64, 70, 72, 80, 84, 90, 96, 100, 112, 120, 126, 128, 140, 144, 160]
    hit = set()
    for P in menu:
        shapes = [(b2c.LINEAR_FULL, P - 9, P - 13, 10, 14),            # full length == P on both axes
                  (b2c.LINEAR_SAME, P - 4, P - 7, 9, 15),              # needs P-4+4, P-7+7
                  (b2c.LINEAR_VALID, P, P - 3, 5, 3),
                  (b2c.CIRCULAR_SAME_PADDED, P - 20, P - 11, 21, 12),
                  (b2c.CIRCULAR_FULL, P // 2, P // 2 + 1, P // 2 + 1, P // 2)]
        for mode, hs, ws, hk, wk in shapes:
            if min(hs, ws, hk, wk) < 1:
                continue
            p = b2c.Plan(mode, hs, ws, hk, wk)
            if (p.h_fft, p.w_fft) != (P, P):
                p.destroy()
                continue
            hit.add(P)
            p.set_path(b2c.PATH_FFT)
            n = 3
            src, ker = rng.random((n, hs, ws)), rng.random((hk, wk))
            d_src, d_ker = torch.from_numpy(src).cuda(), torch.from_numpy(ker).cuda()
            d_dst = torch.full((n, p.h_dst, p.w_dst), float("nan"), dtype=torch.float64, device="cuda")
            p.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), 1, d_dst.data_ptr(), n, 0)
            torch.cuda.synchronize()
            got = d_dst.cpu().numpy()
            for i in range(n):
                _check(got[i], oracle.convolve_direct8(mode, src[i], ker), ("tuned2", P, mode, hs, ws, hk, wk, i))
            g = b2c.Plan(mode, hs, ws, hk, wk)
            g.set_path(b2c.PATH_FFT_GENERIC)
            _check(g.convolve_host(src[1], ker), got[1], ("generic == tuned2", P, mode))
            p.destroy(); g.destroy()
    assert hit == set(menu), sorted(set(menu) - hit)


def test_tuned_square_period_family_with_per_image_kernels(b2c, oracle, torch_cuda):
    """Per-image kernels on every square menu period (16 ... 160, 108 included): each image's kernel runs through the
    forward passes on its own and its spectrum is parked in the CTA's slot (conv_tuned2.cu, PERIMG).  More images
    than resident CTAs would need thousands of them; instead the batch is odd, kernels differ in size up to the
    period itself, and neighbouring images and kernels differ in magnitude by many decades (source and kernel are
    transformed separately, like convolution_fftw.h:190-191, so 1e-12 must hold without any balancing)."""
    torch = torch_cuda
    rng = np.random.default_rng(43)
    menu = [16, 20, 24, 28, 32, 36, 40, 42, 48, 50, 54, 56, 60, 64, 70, 72, 80, 84, 90, 96, 100, 108, 112, 120, 126, 128, 140, 144, 160]
    hit = set()
    for P in menu:
        shapes = [(b2c.LINEAR_SAME, P - 4, P - 7, 9, 15), (b2c.LINEAR_FULL, P - 9, P - 13, 10, 14),
                  (b2c.CIRCULAR_FULL, P // 2, P // 2 + 1, P // 2 + 1, P // 2), (b2c.LINEAR_VALID, P, P - 3, 5, 3)]
        for mode, hs, ws, hk, wk in shapes:
            if min(hs, ws, hk, wk) < 1:
                continue
            p = b2c.Plan(mode, hs, ws, hk, wk)
            if (p.h_fft, p.w_fft) != (P, P):
                p.destroy()
                continue
            hit.add(P)
            p.set_path(b2c.PATH_FFT)
            n = 5
            src = rng.random((n, hs, ws)) * 10.0 ** rng.uniform(-5, 5, size=(n, 1, 1))
            ker = (rng.random((n, hk, wk)) - 0.4) * 10.0 ** rng.uniform(-5, 5, size=(n, 1, 1))
            d_src, d_ker = torch.from_numpy(src).cuda(), torch.from_numpy(ker).cuda()
            d_dst = torch.full((n, p.h_dst, p.w_dst), float("nan"), dtype=torch.float64, device="cuda")
            p.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), 0, d_dst.data_ptr(), n, 0)
            torch.cuda.synchronize()
            got = d_dst.cpu().numpy()
            for i in range(n):
                _check(got[i], oracle.convolve_direct8(mode, src[i], ker[i]), ("tuned2 per-image", P, mode, hs, ws, hk, wk, i))
            p.destroy()
    assert hit == set(menu), sorted(set(menu) - hit)
    # many more images than CTAs: every CTA reuses its spectrum slot many times (headline shape, per-image kernels)
    n, hs, hk = 1500, 100, 15
    src, ker = rng.random((n, hs, hs)), rng.random((n, hk, hk))
    p = b2c.Plan(b2c.LINEAR_SAME, hs, hs, hk, hk)
    d_src, d_ker = torch.from_numpy(src).cuda(), torch.from_numpy(ker).cuda()
    d_dst = torch.full((n, hs, hs), float("nan"), dtype=torch.float64, device="cuda")
    p.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), 0, d_dst.data_ptr(), n, 0)
    torch.cuda.synchronize()
    got = d_dst.cpu().numpy()
    for i in (0, 1, 295, 296, 297, 700, 1498, 1499):
        _check(got[i], oracle.convolve_direct8(b2c.LINEAR_SAME, src[i], ker[i]), ("per-image slots reused", i))
    p.destroy()
