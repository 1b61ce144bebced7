// Drop-in replacement for the reference's Convolution/src/convolution_gsl.h.
//
// The reference's GSL back end has exactly the interface of its FFTW back end (same Workspace
// fields, init_workspace / convolve / clear_workspace, same eight Convolution_Mode enumerators;
// convolution_gsl.h:17-41,43,148,224) and differs only in the CPU FFT library underneath.  On the
// GPU there is one engine, so this header is the same thin wrapper over libb2conv.so in namespace
// GSL_Convolution: a caller built with -DCONVOLUTION=1 compiles and links unchanged.  (Do not
// include it together with convolution_fftw.h in one translation unit unless factorize.h's names
// are acceptable from either: both guard them with FACTORIZE_H.)
#ifndef CONVOLUTION_GSL_H
#define CONVOLUTION_GSL_H

#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "b2conv.h"

#ifndef FACTORIZE_H
#define FACTORIZE_H
// The reference pulls factorize.h in through this header, so find_closest_factor stays visible to
// callers.  Restated: smallest j >= n whose last factor over the given radix list is in the list
// and which is not a multiple of 4 (the `n % 4*4*4*2 == 0` guard of factorize.h:74 as C parses it).
inline bool is_optimal(int n, int* implemented_factors) {
  if ((n % 4) * 32 == 0) return false;
  int rest = n, last = n == 1 ? 1 : 0;
  for (int i = 0; implemented_factors[i] && rest != 1; ++i)
    for (; rest % implemented_factors[i] == 0; rest /= implemented_factors[i]) last = implemented_factors[i];
  if (rest != 1) last = rest;
  for (int i = 0; implemented_factors[i]; ++i)
    if (implemented_factors[i] == last) return true;
  return false;
}
inline int find_closest_factor(int n, int* implemented_factor) {
  int j = n;
  while (!is_optimal(j, implemented_factor)) ++j;
  return j;
}
#endif

namespace GSL_Convolution {

static int GSL_FACTORS[7] = {7, 6, 5, 4, 3, 2, 0};  // kept for source compatibility (reference convolution_gsl.h:15)

typedef enum {
  LINEAR_FULL,
  LINEAR_SAME_UNPADDED,
  LINEAR_SAME,
  LINEAR_VALID,
  CIRCULAR_SAME,
  CIRCULAR_SAME_PADDED,
  CIRCULAR_FULL_UNPADDED,
  CIRCULAR_FULL
} Convolution_Mode;

typedef struct Workspace {
  double *in_src, *out_src, *in_kernel, *out_kernel;  // always NULL here (see header comment)
  int h_src, w_src, h_kernel, w_kernel;
  int w_fftw, h_fftw;
  Convolution_Mode mode;
  double* dst_fft;   // always NULL here
  double* dst;       // the result, h_dst x w_dst row-major, owned by the workspace
  int h_dst, w_dst;  // set by init_workspace
  b2c_plan* plan;    // replaces p_forw_src / p_forw_kernel / p_back
  int dst_pinned;    // dst is page-locked (the device writes it directly); else plain new[]
} Workspace;

inline void init_workspace(Workspace& ws, Convolution_Mode mode, int h_src, int w_src, int h_kernel, int w_kernel) {
  (void)GSL_FACTORS;
  ws.in_src = ws.out_src = ws.in_kernel = ws.out_kernel = ws.dst_fft = NULL;
  ws.h_src = h_src;
  ws.w_src = w_src;
  ws.h_kernel = h_kernel;
  ws.w_kernel = w_kernel;
  ws.mode = mode;
  ws.plan = NULL;
  ws.h_fftw = ws.w_fftw = ws.h_dst = ws.w_dst = 0;
  if (b2c_plan_create(&ws.plan, (int)mode, h_src, w_src, h_kernel, w_kernel) != 0) {
    printf("%s\n", b2c_last_error());
    ws.dst = new double[0];
    ws.dst_pinned = 0;
    return;
  }
  b2c_plan_query(ws.plan, &ws.h_fftw, &ws.w_fftw, &ws.h_dst, &ws.w_dst);
  // page-locked result buffer: the device copies straight into it (12 us less per call than pageable memory)
  void* pinned = NULL;
  ws.dst_pinned = b2c_host_alloc(&pinned, (long)(sizeof(double) * (size_t)ws.h_dst * ws.w_dst)) == 0;
  ws.dst = ws.dst_pinned ? (double*)pinned : new double[(size_t)ws.h_dst * ws.w_dst];
}

inline void clear_workspace(Workspace& ws) {
  if (ws.dst_pinned) b2c_host_free(ws.dst);
  else delete[] ws.dst;
  ws.dst = NULL;
  ws.dst_pinned = 0;
  b2c_plan_destroy(ws.plan);
  ws.plan = NULL;
}

inline void convolve(Workspace& ws, double* src, double* kernel) {
  if (ws.h_fftw <= 0 || ws.w_fftw <= 0) return;
  if (b2c_convolve_host(ws.plan, src, kernel, ws.dst) != 0) printf("%s\n", b2c_last_error());
}

// Additive, not in the reference: convolve `n` images in one call.  src is [n][h_src][w_src],
// kernel [h_kernel][w_kernel] (kernel_is_shared) or [n][h_kernel][w_kernel], dst [n][h_dst][w_dst],
// all caller-owned host memory.
inline void convolve_batch(Workspace& ws, const double* src, const double* kernel, bool kernel_is_shared,
                           double* dst, long n) {
  if (ws.h_fftw <= 0 || ws.w_fftw <= 0) return;
  if (b2c_convolve_batch_host(ws.plan, src, kernel, kernel_is_shared ? 1 : 0, dst, n) != 0)
    printf("%s\n", b2c_last_error());
}

}  // namespace GSL_Convolution

#endif
