// TEST INFRASTRUCTURE ONLY.
// The reference's test driver compiled byte-unchanged with -DCONVOLUTION=1 (its GSL back end) against
// the drop-in include/convolution_gsl.h: same include-guard mechanism as harness_fftw.cc.
#include "../include/convolution_gsl.h"

#include "convolution_test.cc"
