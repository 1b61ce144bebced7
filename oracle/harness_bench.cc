// TEST INFRASTRUCTURE ONLY.
// The reference's benchmark driver, Convolution/src/convolution_benchmark.cc, byte-unchanged, on
// the drop-in header (same include-guard mechanism as harness_fftw.cc): BASELINE configs[0]
// "single 50x50 (*) 10x10 LINEAR_SAME through convolve()" measured exactly the way the reference
// measures it (mean of 20 calls, gettimeofday, planning excluded; :100,169-199).
#include "../include/convolution_fftw.h"

#include "convolution_benchmark.cc"
