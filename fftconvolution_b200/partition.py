"""Batch partitioner and verification helpers.

Images are independent, so N GPUs = N processes that each own a contiguous slice of the batch;
there is no data-path collective.  torch.distributed is used only by callers (bench.py, tests)
to all-reduce the few checksum words these helpers return.
"""
import ctypes

from .capi import check, lib


def partition(n, world, rank):
    """[begin, end) of the n images owned by `rank` of `world`: rank g owns
    [g*ceil(n/world), min(n, (g+1)*ceil(n/world)))."""
    b, e = ctypes.c_long(), ctypes.c_long()
    check(lib().b2c_partition(int(n), int(world), int(rank), ctypes.byref(b), ctypes.byref(e)))
    return b.value, e.value


def checksum_device(ptr, n, stream=0):
    """Order-independent checksum of n device doubles: (wrapping sum of bit patterns, xor of bit
    patterns, plain double sum)."""
    out = (ctypes.c_ulonglong * 3)()
    check(lib().b2c_checksum_device(int(ptr), int(n), out, ctypes.c_void_p(int(stream))))
    import struct
    return out[0], out[1], struct.unpack("<d", struct.pack("<Q", out[2]))[0]


def fill_uniform_device(ptr, n, seed, first=0, stream=0):
    """d[i] = (splitmix64(seed + first + i) >> 11) * 2**-53 on the device."""
    check(lib().b2c_fill_uniform_device(int(ptr), int(n), int(seed), int(first), ctypes.c_void_p(int(stream))))
