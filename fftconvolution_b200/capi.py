"""ctypes binding of include/b2conv.h (the C-ABI of libb2conv.so).

Only plumbing lives here: argument types, error translation.  All arithmetic happens in the CUDA
library; a missing library or a missing GPU is an error, never a fallback.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))

# Convolution_Mode ordinals, reference Convolution/src/convolution_fftw.h:17-27
LINEAR_FULL, LINEAR_SAME_UNPADDED, LINEAR_SAME, LINEAR_VALID = 0, 1, 2, 3
CIRCULAR_SAME, CIRCULAR_SAME_PADDED, CIRCULAR_FULL_UNPADDED, CIRCULAR_FULL = 4, 5, 6, 7
MODE_NAMES = [
    "LINEAR_FULL", "LINEAR_SAME_UNPADDED", "LINEAR_SAME", "LINEAR_VALID",
    "CIRCULAR_SAME", "CIRCULAR_SAME_PADDED", "CIRCULAR_FULL_UNPADDED", "CIRCULAR_FULL",
]
PATH_AUTO, PATH_FFT, PATH_DIRECT, PATH_FFT_GENERIC = 0, 1, 2, 3

# every symbol include/b2conv.h declares -- tests/test_abi.py checks the header against this list
SYMBOLS = [
    "b2c_plan_create", "b2c_plan_query", "b2c_plan_info", "b2c_plan_set_path", "b2c_plan_destroy",
    "b2c_convolve_host", "b2c_convolve_batch_host", "b2c_convolve_batch_device",
    "b2c_convolve_batch_device_static", "b2c_convolve_batch_device_f32", "b2c_convolve_batch_device_c64",
    "b2c_find_closest_factor", "b2c_smooth_size", "b2c_partition", "b2c_checksum_device",
    "b2c_fill_uniform_device", "b2c_multi_convolve_batch_host", "b2c_host_alloc", "b2c_host_free", "b2c_set_device",
    "b2c_launch_count", "b2c_kernel_timer", "b2c_kernel_timer_read", "b2c_last_error", "b2c_version",
]


class B2CError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("b2conv error %d: %s" % (code, message))
        self.code = code


def lib_path():
    # B2C_LIB_PATH: A/B-test another build of the same library (tools/quick_bench.py); never a fallback
    return os.environ.get("B2C_LIB_PATH") or os.path.join(_HERE, "lib", "libb2conv.so")


_lib = None


def lib():
    """Load libb2conv.so (built in-tree by __graft_entry__.build()).  Fails loudly if absent."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise B2CError(-1, "%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback)" % path)
    L = ctypes.CDLL(path)
    c_int, c_long, c_void_p = ctypes.c_int, ctypes.c_long, ctypes.c_void_p
    ip, lp = ctypes.POINTER(c_int), ctypes.POINTER(c_long)
    dp = ctypes.POINTER(ctypes.c_double)
    L.b2c_plan_create.argtypes = [ctypes.POINTER(c_void_p), c_int, c_int, c_int, c_int, c_int]
    L.b2c_plan_query.argtypes = [c_void_p, ip, ip, ip, ip]
    L.b2c_plan_info.argtypes = [c_void_p, ip, ip, ip, ip, ip, lp, ip]
    L.b2c_plan_set_path.argtypes = [c_void_p, c_int]
    L.b2c_plan_destroy.argtypes = [c_void_p]
    L.b2c_convolve_host.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p]
    L.b2c_convolve_batch_host.argtypes = [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_long]
    L.b2c_convolve_batch_device.argtypes = [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_long, c_void_p]
    L.b2c_convolve_batch_device_static.argtypes = [c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                                   c_long, c_void_p]
    L.b2c_convolve_batch_device_f32.argtypes = [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_long, c_void_p]
    L.b2c_convolve_batch_device_c64.argtypes = [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_long, c_void_p]
    L.b2c_find_closest_factor.argtypes = [c_int]
    L.b2c_smooth_size.argtypes = [c_int]
    L.b2c_partition.argtypes = [c_long, c_int, c_int, lp, lp]
    L.b2c_checksum_device.argtypes = [c_void_p, c_long, ctypes.POINTER(ctypes.c_ulonglong), c_void_p]
    L.b2c_fill_uniform_device.argtypes = [c_void_p, c_long, ctypes.c_ulonglong, c_long, c_void_p]
    L.b2c_host_alloc.argtypes = [ctypes.POINTER(c_void_p), c_long]
    L.b2c_host_free.argtypes = [c_void_p]
    L.b2c_set_device.argtypes = [c_int]
    L.b2c_launch_count.restype = c_long
    L.b2c_kernel_timer.argtypes = [c_int]
    L.b2c_kernel_timer_read.argtypes = [dp, lp]
    L.b2c_last_error.restype = ctypes.c_char_p
    L.b2c_version.restype = ctypes.c_char_p
    _lib = L
    return L


def check(rc):
    if rc != 0:
        raise B2CError(rc, lib().b2c_last_error().decode("utf-8", "replace"))
