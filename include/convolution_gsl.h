// Drop-in replacement for the reference's Convolution/src/convolution_gsl.h.
//
// The reference's GSL back end has the interface of its FFTW back end (init_workspace / convolve /
// clear_workspace, the same eight Convolution_Mode enumerators; convolution_gsl.h:17-43,148,224) and differs
// only in the CPU FFT library underneath.  On the GPU there is one engine, so this header is the same thin wrapper
// over libb2conv.so in namespace GSL_Convolution: a caller built with -DCONVOLUTION=1 compiles and links unchanged.
// The Workspace keeps the GSL back end's field names (convolution_gsl.h:30-41): h_fft / w_fft hold the new
// planner's periods, the GSL workspaces, wavetables and matrices are void* placeholders that stay NULL (libgsl is
// not needed to compile against this header).  (Do not include it together with convolution_fftw.h in one
// translation unit unless factorize.h's names are acceptable from either: both guard them with FACTORIZE_H.)
#ifndef CONVOLUTION_GSL_H
#define CONVOLUTION_GSL_H

#define B2C_DROPIN_NAMESPACE GSL_Convolution
#define B2C_DROPIN_FACTORS static int GSL_FACTORS[7] = {7, 6, 5, 4, 3, 2, 0};   /* reference convolution_gsl.h:15 */
#define B2C_DROPIN_LEGACY_FIELDS                                          \
  void *ws_column, *ws_line, *wv_column, *wv_line; /* always NULL here */ \
  int h_fft, w_fft;                                                       \
  void *fft, *fft_copy; /* always NULL here */
#define B2C_DROPIN_LEGACY_INIT(ws) \
  (void)GSL_FACTORS;               \
  ws.ws_column = ws.ws_line = ws.wv_column = ws.wv_line = ws.fft = ws.fft_copy = NULL;
#define B2C_DROPIN_HFFT h_fft
#define B2C_DROPIN_WFFT w_fft
#include "b2conv_dropin.inc"

#endif
