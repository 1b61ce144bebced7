"""Test-side loader for the CPU oracle (oracle/_build/liboracle.so) and, where it was built, the
unmodified reference behind extern "C" (oracle/_ref/libref.so).  TEST INFRASTRUCTURE ONLY."""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "_build", "liboracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libref.so")

MODES = ["LINEAR_FULL", "LINEAR_SAME_UNPADDED", "LINEAR_SAME", "LINEAR_VALID", "CIRCULAR_SAME",
         "CIRCULAR_SAME_PADDED", "CIRCULAR_FULL_UNPADDED", "CIRCULAR_FULL"]


def _ptr(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


class _Lib:
    def __init__(self, path, prefix):
        self.L = ctypes.CDLL(path)
        self.prefix = prefix

    def plan(self, mode, hs, ws, hk, wk):
        v = [ctypes.c_int() for _ in range(4)]
        fn = getattr(self.L, "orc_plan" if self.prefix == "orc" else "ref_fftw_plan")
        fn(mode, hs, ws, hk, wk, *[ctypes.byref(x) for x in v])
        return tuple(x.value for x in v)  # h_fft, w_fft, h_dst, w_dst

    def find_closest_factor(self, n):
        return getattr(self.L, self.prefix + "_find_closest_factor")(n)

    def _run(self, name, mode, src, ker, shape):
        src = np.ascontiguousarray(src, dtype=np.float64)
        ker = np.ascontiguousarray(ker, dtype=np.float64)
        dst = np.zeros(shape if shape[0] * shape[1] > 0 else (0, 0), dtype=np.float64)
        buf = dst if dst.size else np.zeros(1)
        getattr(self.L, name)(mode, src.shape[0], src.shape[1], ker.shape[0], ker.shape[1], _ptr(src), _ptr(ker), _ptr(buf))
        return dst

    def convolve_fftw(self, mode, src, ker):
        """The reference FFT path: restated (oracle) or the unmodified header over the shim (ref)."""
        _, _, hd, wd = self.plan(mode, src.shape[0], src.shape[1], ker.shape[0], ker.shape[1])
        name = "orc_convolve_fftw" if self.prefix == "orc" else "ref_fftw_convolve"
        return self._run(name, mode, src, ker, (hd, wd))

    def convolve_direct8(self, mode, src, ker):
        """All eight modes from nested loops with long-double accumulation (oracle only)."""
        _, _, hd, wd = self.plan(mode, src.shape[0], src.shape[1], ker.shape[0], ker.shape[1])
        return self._run("orc_convolve_direct8", mode, src, ker, (hd, wd))

    def convolve_std(self, std_mode, src, ker):
        hs, ws = src.shape
        hk, wk = ker.shape
        shp = {0: (hs + hk - 1, ws + wk - 1), 1: (hs, ws), 2: (hs - hk + 1, ws - wk + 1), 3: (hs, ws),
               4: (hs + hk - 1, ws + wk - 1)}[std_mode]
        if shp[0] <= 0 or shp[1] <= 0:
            return np.zeros((0, 0))
        name = "orc_convolve_std" if self.prefix == "orc" else "ref_std_convolve"
        return self._run(name, std_mode, src, ker, shp)

    def batch_fftw(self, mode, src, ker, threads=1):
        n, hs, ws = src.shape
        shared = 1 if ker.ndim == 2 else 0
        hk, wk = ker.shape[-2:]
        _, _, hd, wd = self.plan(mode, hs, ws, hk, wk)
        dst = np.zeros((n, hd, wd))
        name = "orc_convolve_fftw_batch" if self.prefix == "orc" else "ref_fftw_batch"
        fn = getattr(self.L, name)
        fn.argtypes = [ctypes.c_int] * 5 + [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                                            ctypes.c_long, ctypes.c_int]
        src = np.ascontiguousarray(src)
        ker = np.ascontiguousarray(ker)
        fn(mode, hs, ws, hk, wk, src.ctypes.data, ker.ctypes.data, shared, dst.ctypes.data if dst.size else None, n, threads)
        return dst


    def batch_std(self, std_mode, src, ker, threads=1):
        """The reference's nested loops (convolution_std.h) over a batch; unmodified-reference library only."""
        assert self.prefix == "ref"
        n, hs, ws = src.shape
        hk, wk = ker.shape[-2:]
        shp = {0: (hs + hk - 1, ws + wk - 1), 1: (hs, ws), 2: (hs - hk + 1, ws - wk + 1), 3: (hs, ws),
               4: (hs + hk - 1, ws + wk - 1)}[std_mode]
        dst = np.zeros((n,) + shp)
        fn = self.L.ref_std_batch
        fn.argtypes = [ctypes.c_int] * 5 + [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                                            ctypes.c_long, ctypes.c_int]
        src, ker = np.ascontiguousarray(src), np.ascontiguousarray(ker)
        fn(std_mode, hs, ws, hk, wk, src.ctypes.data, ker.ctypes.data, 1 if ker.ndim == 2 else 0, dst.ctypes.data, n, threads)
        return dst


def build_oracle():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle")])


def load_oracle():
    if not os.path.exists(ORACLE_SO):
        build_oracle()
    return _Lib(ORACLE_SO, "orc")


def load_ref():
    """The unmodified reference (None where oracle/_ref was not built / did not travel)."""
    if not os.path.exists(REF_SO):
        return None
    return _Lib(REF_SO, "ref")


def rel_err(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    if b.size == 0:
        return 0.0 if a.size == 0 else float("inf")
    den = np.linalg.norm(b.ravel())
    return float(np.linalg.norm((a - b).ravel()) / (den if den > 0 else 1.0))


# counter-based generator shared with the device (util_kernels.cu): splitmix64(seed + i) >> 11 * 2^-53
def splitmix_uniform(seed, first, n):
    with np.errstate(over="ignore"):
        x = (np.arange(first, first + n, dtype=np.uint64) + np.uint64(seed)) + np.uint64(0x9E3779B97F4A7C15)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        x = x ^ (x >> np.uint64(31))
    return (x >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)
