"""fftconvolution_b200 -- B200-native batched FP64 convolution behind the FFTConvolution header API.

The product is ``lib/libb2conv.so`` (hand-written sm_100a CUDA behind the C-ABI of
``include/b2conv.h``).  This package is the Python host-side mirror of the reference's operator
interface (``Workspace`` / ``init_workspace`` / ``convolve`` / ``clear_workspace`` and the
``Convolution_Mode`` ordinals of Convolution/src/convolution_fftw.h:17-42) plus the batch
partitioner.  There is no CPU fallback: without the built library every compute call raises.
"""
from .capi import (  # noqa: F401
    B2CError,
    lib,
    lib_path,
    LINEAR_FULL,
    LINEAR_SAME_UNPADDED,
    LINEAR_SAME,
    LINEAR_VALID,
    CIRCULAR_SAME,
    CIRCULAR_SAME_PADDED,
    CIRCULAR_FULL_UNPADDED,
    CIRCULAR_FULL,
    MODE_NAMES,
    PATH_AUTO,
    PATH_FFT,
    PATH_DIRECT,
    PATH_FFT_GENERIC,
)
from .convolution import (  # noqa: F401
    Plan,
    Workspace,
    init_workspace,
    convolve,
    clear_workspace,
    find_closest_factor,
    smooth_size,
    multi_convolve_batch_host,
)
from .partition import partition, checksum_device, fill_uniform_device  # noqa: F401
