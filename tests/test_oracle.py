"""CPU tests: pin the oracle restatement against (a) the committed golden vectors, which were
produced by the UNMODIFIED reference headers (tests/golden/make_golden.py), and (b) the unmodified
reference itself when oracle/_ref was built on this machine."""
import os

import numpy as np
import pytest

import _oracle

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "golden.npz"))
CASES = [tuple(int(v) for v in row) for row in GOLD["cases"]]
TOL = 1e-12  # BASELINE.json: <= 1e-12 relative vs the reference FFTW path


def _inputs(case):
    hs, ws, hk, wk, seed = case
    rng = np.random.default_rng(seed)
    return rng.random((hs, ws)), rng.random((hk, wk))


def test_planner_matches_reference_table(oracle):
    fcf = GOLD["fcf"]
    got = np.array([oracle.find_closest_factor(n) for n in range(len(fcf))])
    assert np.array_equal(got, fcf)
    # values SURVEY.md 8a-5 quotes for factorize.h:92-104 (multiples of 4 are never returned)
    for n, want in [(0, 2), (1, 2), (2, 2), (4, 5), (8, 9), (100, 105), (107, 110), (108, 110), (128, 130),
                    (199, 210), (200, 210), (4096, 4125), (8191, 8250)]:
        assert oracle.find_closest_factor(n) == want
    assert not any(v % 4 == 0 for v in got)


@pytest.mark.parametrize("ci", range(len(CASES)))
def test_oracle_fft_path_matches_golden(oracle, ci):
    src, ker = _inputs(CASES[ci])
    for mode in range(8):
        key = "c%d_m%d" % (ci, mode)
        if key not in GOLD:
            continue
        want = GOLD[key]
        got = oracle.convolve_fftw(mode, src, ker)
        assert got.shape == want.shape, (mode, got.shape, want.shape)
        assert _oracle.rel_err(got, want) <= TOL, (mode, CASES[ci])
        # the nested-loop restatement (all eight modes through the identities of SURVEY 8a-8)
        got2 = oracle.convolve_direct8(mode, src, ker)
        assert got2.shape == want.shape
        assert _oracle.rel_err(got2, want) <= TOL, (mode, CASES[ci])


@pytest.mark.parametrize("ci", range(len(CASES)))
def test_oracle_nested_loops_match_golden(oracle, ci):
    src, ker = _inputs(CASES[ci])
    for sm in range(5):
        key = "c%d_s%d" % (ci, sm)
        if key not in GOLD:
            continue
        want = GOLD[key]
        got = oracle.convolve_std(sm, src, ker)
        assert got.shape == want.shape
        if want.size:
            assert np.array_equal(got, want), "convolution_std.h restatement must be bit-exact"


def test_oracle_shapes_match_reference_rules(oracle):
    # init_workspace table, convolution_fftw.h:52-132
    assert oracle.plan(2, 100, 100, 15, 15) == (110, 110, 100, 100)
    assert oracle.plan(2, 50, 50, 10, 10) == (55, 55, 50, 50)
    assert oracle.plan(7, 100, 100, 100, 100) == (199, 199, 199, 199)
    assert oracle.plan(3, 10, 10, 11, 3) == (0, 0, 0, 0)
    assert oracle.plan(0, 1, 60, 1, 20) == (2, 81, 1, 79)
    assert oracle.plan(4, 1, 60, 1, 20) == (1, 60, 1, 60)


def test_oracle_against_unmodified_reference_here(oracle):
    ref = _oracle.load_ref()
    if ref is None:
        pytest.skip("oracle/_ref not built on this machine (reference sources absent)")
    rng = np.random.default_rng(7)
    for _ in range(25):
        hs, ws, hk, wk = (int(v) for v in rng.integers(1, 40, size=4))
        src, ker = rng.random((hs, ws)), rng.random((hk, wk))
        for mode in range(8):
            assert oracle.plan(mode, hs, ws, hk, wk) == ref.plan(mode, hs, ws, hk, wk)
            a, b = oracle.convolve_fftw(mode, src, ker), ref.convolve_fftw(mode, src, ker)
            assert a.shape == b.shape
            assert _oracle.rel_err(a, b) <= TOL


def test_splitmix_generator_known_answers():
    # splitmix64 reference outputs for state 0 (first three of the published test vector)
    with np.errstate(over="ignore"):
        def sm(x):
            x = (np.uint64(x) + np.uint64(0x9E3779B97F4A7C15))
            x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
            x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
            return int(x ^ (x >> np.uint64(31)))
        assert sm(0) == 0xE220A8397B1DCDAF
    u = _oracle.splitmix_uniform(0x5EED0002, 0, 1000)
    assert u.min() >= 0.0 and u.max() < 1.0 and abs(u.mean() - 0.5) < 0.05
