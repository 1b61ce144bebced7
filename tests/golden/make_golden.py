"""Generate tests/golden/golden.npz by running the UNMODIFIED reference in this container.

  * FFT path   : Convolution/src/convolution_fftw.h, compiled as is over oracle/shim/fftw3.h
                 (no libfftw3 exists here) -> oracle/_ref/libref.so : ref_fftw_convolve
  * nested loops: Convolution/src/convolution_std.h, compiled as is -> ref_std_convolve

Run from the repo root where /root/reference exists:  python tests/golden/make_golden.py
The inputs are regenerated from the stored seeds (numpy default_rng) so the file stays small.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import _oracle  # noqa: E402

CASES = [
    # (hs, ws, hk, wk)  -- BASELINE configs, the reference test categories (convolution_test.cc:424-523), edges
    (50, 50, 10, 10),      # config 1
    (100, 100, 15, 15),    # config 2 shape
    (33, 47, 21, 39),      # 2D category, rectangular
    (40, 30, 20, 29),
    (60, 1, 98, 1),        # 1D even/even Hx1, kernel longer than source (LINEAR_VALID empty)
    (98, 1, 41, 1),        # even/odd
    (61, 1, 20, 1),        # odd/even
    (99, 1, 39, 1),        # odd/odd
    (1, 60, 1, 58),        # 1xW even/even
    (1, 98, 1, 41),
    (1, 61, 1, 20),
    (1, 99, 1, 39),
    (20, 20, 45, 31),      # kernel larger than source on both axes
    (7, 9, 7, 9),          # kernel == source
    (1, 1, 1, 1),
    (2, 3, 1, 1),
    (5, 5, 3, 3),
    (16, 16, 4, 4),
    (1, 64, 1, 65),
    (31, 37, 8, 8),
]


def main():
    ref = _oracle.load_ref()
    if ref is None:
        raise SystemExit("oracle/_ref/libref.so missing: run `make -C oracle` where /root/reference exists")
    out = {}
    meta = []
    for ci, (hs, ws, hk, wk) in enumerate(CASES):
        rng = np.random.default_rng(1000 + ci)
        src = rng.random((hs, ws))
        ker = rng.random((hk, wk))
        big = hs * ws >= 2500  # keep the fixture small: the two BASELINE shapes store fewer modes
        for mode in ((2, 4) if big else range(8)):
            d = ref.convolve_fftw(mode, src, ker)
            out["c%d_m%d" % (ci, mode)] = d
        for sm in ((1,) if big else range(5)):
            out["c%d_s%d" % (ci, sm)] = ref.convolve_std(sm, src, ker)
        meta.append((hs, ws, hk, wk, 1000 + ci))
    out["cases"] = np.array(meta, dtype=np.int64)
    # planner table: find_closest_factor for 0..8200 (factorize.h:92-104)
    out["fcf"] = np.array([ref.find_closest_factor(n) for n in range(0, 8201)], dtype=np.int32)
    path = os.path.join(HERE, "golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
