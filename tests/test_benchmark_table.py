"""the GPU timings in the format of the reference's published tables
(Convolution/Data/benchmarks_convolution_*.txt, written by convolution_benchmark.cc:199 as `mode src kernel seconds`).
The committed table (profiles/r02_benchmarks_convolution_b2c.txt, tools/emit_benchmark_table.py on one B200) must load
the way the reference's plot script loads its own tables (PythonScripts/plot_comparisons.py:61-105: genfromtxt as
strings, rows selected by mode name, a (kernel - min_kernel, src - min_src) grid) and cover the reference's whole grid:
src 4..100, kernel 3..src-1, four mode names (run_fftw_bench.sh:13-45, convolution_benchmark.cc:70-97).
The script itself is Python 2 (print statements, np.float) and cannot run here; its loader is restated below."""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TABLE = os.path.join(ROOT, "profiles", "r02_benchmarks_convolution_b2c.txt")


def _load_like_plot_comparisons(path, mode):
    data = np.genfromtxt(path, dtype="str")
    rows = np.array(data[np.where(data[:, 0] == mode)][:, 1:], dtype=float)
    min_src, max_src = rows[:, 0].min(), rows[:, 0].max()
    min_kernel, max_kernel = rows[:, 1].min(), rows[:, 1].max()
    X, _ = np.meshgrid(np.arange(min_src, max_src + 1), np.arange(min_kernel, max_kernel + 1))
    Z = np.nan * np.zeros(X.shape)
    for d in rows:
        Z[int(d[1] - min_kernel), int(d[0] - min_src)] = d[2]
    return (min_src, max_src, min_kernel, max_kernel), Z


def test_table_loads_like_the_reference_tables_and_covers_the_grid():
    for mode in ("linear", "circular", "linear_unpadded", "circular_unpadded"):
        (s0, s1, k0, k1), Z = _load_like_plot_comparisons(TABLE, mode)
        assert (s0, s1, k0, k1) == (4, 100, 3, 99), (mode, s0, s1, k0, k1)
        n = 0
        for s in range(4, 101):
            for k in range(3, s):
                v = Z[k - 3, s - 4]
                assert np.isfinite(v) and 0 < v < 1e-3, (mode, s, k, v)   # seconds per convolution, batch-amortised
                n += 1
        assert n == 4753 and np.isfinite(Z).sum() == n, mode
