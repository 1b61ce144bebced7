// Drop-in replacement for the reference's Convolution/src/convolution_fftw.h.
//
// Same namespace (FFTW_Convolution), same Convolution_Mode enumerators in the same order
// (reference :17-27), same Workspace field names (:29-42), same three functions with the same
// signatures (:44, :150, :216) -- but the work is done by libb2conv.so (hand-written sm_100a CUDA,
// include/b2conv.h) instead of libfftw3.  A caller such as the reference's own
// convolution_test.cc / convolution_benchmark.cc compiles against this header unchanged and links
// with -lb2conv instead of `pkg-config --libs fftw3`.
//
// What differs, on purpose:
//   * ws.h_fftw / ws.w_fftw hold the NEW planner's 2/3/5/7-smooth periods (no caller reads them);
//     ws.h_dst / ws.w_dst are bit-for-bit the reference's.
//   * in_src / out_src / in_kernel / out_kernel / dst_fft stay NULL: the padded signals, spectra and
//     the uncropped result only ever exist in GPU shared memory.
//   * errors: the reference prints and carries on (:122-131, :273-282); so does this header,
//     printing b2c_last_error().  There is no CPU fallback.
//   * the functions are `inline`, so unlike the reference the header may be included in more
//     than one translation unit.
#ifndef CONVOLUTION_FFTW_H
#define CONVOLUTION_FFTW_H

#define B2C_DROPIN_NAMESPACE FFTW_Convolution
#define B2C_DROPIN_FACTORS static int FFTW_FACTORS[7] = {13, 11, 7, 5, 3, 2, 0};   /* reference :15 */
#define B2C_DROPIN_LEGACY_FIELDS                                                      \
  double *in_src, *out_src, *in_kernel, *out_kernel; /* always NULL here */            \
  int w_fftw, h_fftw;                                                                 \
  double* dst_fft; /* always NULL here */
#define B2C_DROPIN_LEGACY_INIT(ws) \
  (void)FFTW_FACTORS;              \
  ws.in_src = ws.out_src = ws.in_kernel = ws.out_kernel = ws.dst_fft = NULL;
#define B2C_DROPIN_HFFT h_fftw
#define B2C_DROPIN_WFFT w_fftw
#include "b2conv_dropin.inc"

#endif
