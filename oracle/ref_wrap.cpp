// TEST INFRASTRUCTURE ONLY.
// Thin extern "C" wrapper around the UNMODIFIED reference headers, included from where they lie
// under /root/reference (passed as -I by oracle/Makefile; nothing is copied into this repo):
//   Convolution/src/convolution_fftw.h  (over the stand-in oracle/shim/fftw3.h -- no libfftw3 here)
//   Convolution/src/convolution_std.h
//   Convolution/src/factorize.h
// Output: oracle/_ref/libref.so (git-ignored).  Used to pin the oracle restatement, to generate
// tests/golden/, and as the "reference" CPU baseline in bench.py.
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

#include "convolution_fftw.h"
#include "convolution_std.h"

extern "C" {

int ref_find_closest_factor(int n) { return find_closest_factor(n, FFTW_Convolution::FFTW_FACTORS); }

int ref_fftw_plan(int mode, int hs, int ws, int hk, int wk, int* h_fft, int* w_fft, int* h_dst, int* w_dst) {
  FFTW_Convolution::Workspace w;
  FFTW_Convolution::init_workspace(w, (FFTW_Convolution::Convolution_Mode)mode, hs, ws, hk, wk);
  *h_fft = w.h_fftw; *w_fft = w.w_fftw; *h_dst = w.h_dst; *w_dst = w.w_dst;
  FFTW_Convolution::clear_workspace(w);
  return 0;
}

int ref_fftw_convolve(int mode, int hs, int ws, int hk, int wk, const double* src, const double* kernel,
                      double* dst) {
  FFTW_Convolution::Workspace w;
  FFTW_Convolution::init_workspace(w, (FFTW_Convolution::Convolution_Mode)mode, hs, ws, hk, wk);
  FFTW_Convolution::convolve(w, const_cast<double*>(src), const_cast<double*>(kernel));
  if (w.h_fftw > 0 && w.w_fftw > 0)
    std::memcpy(dst, w.dst, sizeof(double) * (size_t)w.h_dst * w.w_dst);
  FFTW_Convolution::clear_workspace(w);
  return 0;
}

int ref_std_convolve(int mode, int hs, int ws, int hk, int wk, const double* src, const double* kernel,
                     double* dst) {
  STD_Convolution::Workspace w;
  STD_Convolution::init_workspace(w, (STD_Convolution::Convolution_Mode)mode, hs, ws, hk, wk);
  STD_Convolution::convolve(w, const_cast<double*>(src), const_cast<double*>(kernel));
  if (w.h_dst > 0 && w.w_dst > 0) {
    std::memcpy(dst, w.dst, sizeof(double) * (size_t)w.h_dst * w.w_dst);
    STD_Convolution::clear_workspace(w);
  }
  return 0;
}

// Batch drivers for the CPU baseline: `threads` workers, each with its own Workspace initialised
// once (plan creation excluded from nothing -- the caller times the whole call; it is amortised
// over the worker's share of the batch, as in convolution_benchmark.cc:169-186).
int ref_fftw_batch(int mode, int hs, int ws, int hk, int wk, const double* src, const double* kernel,
                   int kernel_is_shared, double* dst, long n, int threads) {
  if (threads < 1) threads = 1;
  std::vector<FFTW_Convolution::Workspace> wsp(threads);
  for (int t = 0; t < threads; ++t)  // serialised: the reference's planner is not re-entrant
    FFTW_Convolution::init_workspace(wsp[t], (FFTW_Convolution::Convolution_Mode)mode, hs, ws, hk, wk);
  const size_t dn = (size_t)wsp[0].h_dst * wsp[0].w_dst;
  auto work = [&](int t) {
    FFTW_Convolution::Workspace& w = wsp[t];
    for (long i = t; i < n; i += threads) {
      FFTW_Convolution::convolve(w, const_cast<double*>(src + (size_t)i * hs * ws),
                                 const_cast<double*>(kernel + (kernel_is_shared ? 0 : (size_t)i * hk * wk)));
      if (dst && dn) std::memcpy(dst + (size_t)i * dn, w.dst, sizeof(double) * dn);
    }
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < threads; ++t) pool.emplace_back(work, t);
  work(0);
  for (auto& th : pool) th.join();
  for (int t = 0; t < threads; ++t) FFTW_Convolution::clear_workspace(wsp[t]);
  return 0;
}

int ref_std_batch(int mode, int hs, int ws, int hk, int wk, const double* src, const double* kernel,
                  int kernel_is_shared, double* dst, long n, int threads) {
  if (threads < 1) threads = 1;
  std::vector<STD_Convolution::Workspace> wsp(threads);
  for (int t = 0; t < threads; ++t)
    STD_Convolution::init_workspace(wsp[t], (STD_Convolution::Convolution_Mode)mode, hs, ws, hk, wk);
  if (wsp[0].h_dst <= 0 || wsp[0].w_dst <= 0) return 0;
  const size_t dn = (size_t)wsp[0].h_dst * wsp[0].w_dst;
  auto work = [&](int t) {
    STD_Convolution::Workspace& w = wsp[t];
    for (long i = t; i < n; i += threads) {
      STD_Convolution::convolve(w, const_cast<double*>(src + (size_t)i * hs * ws),
                                const_cast<double*>(kernel + (kernel_is_shared ? 0 : (size_t)i * hk * wk)));
      if (dst) std::memcpy(dst + (size_t)i * dn, w.dst, sizeof(double) * dn);
    }
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < threads; ++t) pool.emplace_back(work, t);
  work(0);
  for (auto& th : pool) th.join();
  for (int t = 0; t < threads; ++t) STD_Convolution::clear_workspace(wsp[t]);
  return 0;
}

}  // extern "C"
