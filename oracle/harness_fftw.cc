// TEST INFRASTRUCTURE ONLY.
// Acceptance harness: the reference's own test driver, Convolution/src/convolution_test.cc,
// compiled BYTE-UNCHANGED (it is #included from where it lies under /root/reference, nothing is
// copied) against the drop-in header include/convolution_fftw.h of this repo.
//
// How the unchanged driver ends up on the new header: this file includes include/convolution_fftw.h
// first; it uses the same include guard as the reference header (CONVOLUTION_FFTW_H), so the
// driver's own `#include "convolution_fftw.h"` -- which resolves to the reference file next to it --
// expands to nothing.  Octave is replaced by oracle/octave_shim (CPU oracle).
// Built by oracle/Makefile with -DCONVOLUTION=0 into oracle/_ref/convolution_test_b2c.bin.
#include "../include/convolution_fftw.h"

#include "convolution_test.cc"
