"""CPU tests of the boundary: the library loads without a GPU, exports every symbol the header
declares, and its host-side planning reproduces the reference's output shapes bit for bit."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "b2conv.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b2c_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(b2c):
    from fftconvolution_b200 import capi
    L = b2c.lib()
    declared = _header_symbols()
    assert declared, "no declarations parsed from include/b2conv.h"
    for name in declared:
        assert hasattr(L, name), "libb2conv.so does not export %s" % name
    assert sorted(capi.SYMBOLS) == declared
    assert b"sm_100a" in L.b2c_version()


def test_dropin_headers_compile_with_the_reference_names(tmp_path):
    """Both drop-in headers (one shared body, include/b2conv_dropin.inc) compile as plain C++ and expose the names a
    caller of the reference uses: namespace, enum ordinals, functions, the factor list, and each back end's own
    Workspace field names (convolution_fftw.h:29-42, convolution_gsl.h:30-41)."""
    import subprocess
    src = tmp_path / "t.cc"
    src.write_text("""
#include "convolution_fftw.h"
#include "convolution_gsl.h"
static_assert(FFTW_Convolution::LINEAR_FULL == 0 && FFTW_Convolution::LINEAR_SAME_UNPADDED == 1 &&
              FFTW_Convolution::LINEAR_SAME == 2 && FFTW_Convolution::LINEAR_VALID == 3 &&
              FFTW_Convolution::CIRCULAR_SAME == 4 && FFTW_Convolution::CIRCULAR_SAME_PADDED == 5 &&
              FFTW_Convolution::CIRCULAR_FULL_UNPADDED == 6 && FFTW_Convolution::CIRCULAR_FULL == 7, "ordinals");
static_assert(GSL_Convolution::CIRCULAR_FULL == 7, "ordinals");
void (*f1)(FFTW_Convolution::Workspace&, FFTW_Convolution::Convolution_Mode, int, int, int, int) = FFTW_Convolution::init_workspace;
void (*f2)(FFTW_Convolution::Workspace&, double*, double*) = FFTW_Convolution::convolve;
void (*f3)(FFTW_Convolution::Workspace&) = FFTW_Convolution::clear_workspace;
void (*g1)(GSL_Convolution::Workspace&, GSL_Convolution::Convolution_Mode, int, int, int, int) = GSL_Convolution::init_workspace;
void (*g2)(GSL_Convolution::Workspace&, double*, double*) = GSL_Convolution::convolve;
void (*g3)(GSL_Convolution::Workspace&) = GSL_Convolution::clear_workspace;
int use(FFTW_Convolution::Workspace& a, GSL_Convolution::Workspace& b) {
  a.in_src = a.out_src = a.in_kernel = a.out_kernel = a.dst_fft = 0;
  b.ws_column = b.ws_line = b.wv_column = b.wv_line = b.fft = b.fft_copy = 0;
  return a.h_fftw + a.w_fftw + a.h_dst + a.w_dst + b.h_fft + b.w_fft + b.h_dst + b.w_dst +
         find_closest_factor(107, FFTW_Convolution::FFTW_FACTORS) + GSL_Convolution::GSL_FACTORS[0] + (a.dst == b.dst);
}
""")
    subprocess.check_call(["g++", "-std=c++17", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-c", str(src),
                           "-o", str(tmp_path / "t.o")])


def test_mode_ordinals(b2c):
    assert [b2c.LINEAR_FULL, b2c.LINEAR_SAME_UNPADDED, b2c.LINEAR_SAME, b2c.LINEAR_VALID, b2c.CIRCULAR_SAME,
            b2c.CIRCULAR_SAME_PADDED, b2c.CIRCULAR_FULL_UNPADDED, b2c.CIRCULAR_FULL] == list(range(8))


def test_output_shapes_match_reference_for_sweep_grid(b2c, oracle):
    """config 5 grid: source 1..100 x kernel 1..100, all eight modes (square), plus rectangles."""
    bad = []
    for mode in range(8):
        for hs in list(range(1, 101, 3)) + [100]:
            for hk in list(range(1, 101, 3)) + [100]:
                p = b2c.Plan(mode, hs, hs, hk, hk)
                _, _, hd, wd = oracle.plan(mode, hs, hs, hk, hk)
                if (p.h_dst, p.w_dst) != (hd, wd):
                    bad.append((mode, hs, hk))
                if hd > 0:
                    assert p.h_fft >= 1 and p.w_fft >= 1
                else:
                    assert p.h_fft == 0 and p.w_fft == 0
                p.destroy()
    assert not bad, bad[:5]
    rng = np.random.default_rng(3)
    for _ in range(300):
        hs, ws, hk, wk = (int(v) for v in rng.integers(1, 120, size=4))
        mode = int(rng.integers(0, 8))
        p = b2c.Plan(mode, hs, ws, hk, wk)
        assert (p.h_dst, p.w_dst) == oracle.plan(mode, hs, ws, hk, wk)[2:]
        p.destroy()


def test_periods_are_smooth_and_alias_safe(b2c):
    def smooth(n):
        for q in (2, 3, 5, 7):
            while n % q == 0:
                n //= q
        return n == 1
    for mode in range(8):
        for hs, hk in [(100, 15), (97, 15), (50, 10), (30, 45), (1, 1), (64, 64), (4096, 4097), (4095, 3)]:
            p = b2c.Plan(mode, 1, hs, 1, hk)
            if p.w_dst == 0:
                continue
            assert smooth(p.w_fft) and p.h_fft == 1
            i = p.info()
            L = hs + hk - 1
            if mode == b2c.CIRCULAR_SAME:
                assert p.w_fft == hs or (i["fold_w"] == hs and p.w_fft >= hs + min(hs, hk) - 1)
            else:
                off, n = i["off_w"], p.w_dst
                assert p.w_fft >= max(L - off, off + n)   # no aliasing inside the kept window
            p.destroy()
    p = b2c.Plan(b2c.LINEAR_SAME, 100, 100, 15, 15)
    assert p.h_fft >= 107 and p.w_fft >= 107
    assert b2c.smooth_size(107) >= 107


def test_reference_planner_restated_in_library(b2c, oracle):
    for n in list(range(0, 600)) + [4096, 4564, 8191]:
        assert b2c.find_closest_factor(n) == oracle.find_closest_factor(n)


def test_unknown_mode_is_an_error_with_the_mode_list(b2c):
    with pytest.raises(b2c.B2CError) as e:
        b2c.Plan(8, 4, 4, 2, 2)
    assert "LINEAR_FULL" in str(e.value) and "CIRCULAR_FULL" in str(e.value)
    ws = b2c.Workspace()
    b2c.init_workspace(ws, 11, 4, 4, 2, 2)   # prints and carries on, like convolution_fftw.h:122-131
    assert ws.plan is None
    b2c.convolve(ws, np.zeros((4, 4)), np.zeros((2, 2)))  # no-op


def test_empty_valid_plan_is_a_noop_without_gpu(b2c):
    ws = b2c.Workspace()
    b2c.init_workspace(ws, b2c.LINEAR_VALID, 10, 10, 11, 3)
    assert (ws.h_dst, ws.w_dst, ws.h_fftw, ws.w_fftw) == (0, 0, 0, 0)
    b2c.convolve(ws, np.zeros((10, 10)), np.zeros((11, 3)))
    b2c.clear_workspace(ws)


def test_partition_covers_batch(b2c):
    for n in [0, 1, 7, 64, 65536, 65537]:
        for world in [1, 2, 4, 8]:
            spans = [b2c.partition(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            for a, b in zip(spans, spans[1:]):
                assert a[1] == b[0]
            per = -(-n // world) if n else 0
            assert all(e - b <= per for b, e in spans)
    with pytest.raises(b2c.B2CError):
        b2c.partition(10, 2, 2)


def test_size_limits_of_planning(b2c):
    """Planning is host-only: transforms past the any-size kernel's index range are accepted when strip
    mode can run them (no folding, period <= 4800 rows) and refuse the generic path; others fail with
    B2C_ESIZE and a message."""
    p = b2c.Plan(b2c.LINEAR_SAME, 4096, 4096, 15, 15)
    assert (p.h_dst, p.w_dst) == (4096, 4096) and p.h_fft >= 4103
    with pytest.raises(b2c.B2CError) as e:
        p.set_path(b2c.PATH_FFT_GENERIC)
    assert e.value.code == 4 and "strip mode" in str(e.value)
    p.destroy()
    with pytest.raises(b2c.B2CError) as e:
        b2c.Plan(b2c.CIRCULAR_SAME, 4099, 4099, 3, 3)      # prime size: fold path only
    assert e.value.code == 4
    with pytest.raises(b2c.B2CError) as e:
        b2c.Plan(b2c.LINEAR_FULL, 6000, 6000, 3, 3)
    assert e.value.code == 4
