// TEST INFRASTRUCTURE ONLY.
// CPU restatement ("oracle") of the reference hot path of jeremyfix/FFTConvolution:
//   * size planner            -> Convolution/src/factorize.h:7-104
//   * init_workspace shapes   -> Convolution/src/convolution_fftw.h:44-132
//   * fftw_circular_convolution + convolve -> Convolution/src/convolution_fftw.h:169-284
//   * nested-loop convolution -> Convolution/src/convolution_std.h:86-199
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library, and only as the checker.  The product (libb2conv.so) never links it.
//
// Parity pin: there are no golden vectors in the reference tree and neither libfftw3 nor Octave
// exist in the build container, so this restatement is pinned against outputs of the UNMODIFIED
// reference headers run here (oracle/_ref: convolution_std.h as is, convolution_fftw.h over the
// stand-in fftw3.h in oracle/shim/) -- see tests/golden/make_golden.py and tests/test_oracle.py.
//
// Build: oracle/Makefile -> oracle/_build/liboracle.so
#include <algorithm>
#include <cmath>
#include <complex>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

#include "mixed_fft.h"

namespace {

// ---------------------------------------------------------------- planner (factorize.h)
const int kFactors[7] = {13, 11, 7, 5, 3, 2, 0};  // FFTW_FACTORS, convolution_fftw.h:15

// factorize.h:71-90.  NOTE the guard `n % 4*4*4*2 == 0` parses as ((n%4)*32)==0, i.e. every
// multiple of 4 is rejected; restated literally.
bool is_optimal(int n) {
  if (n % 4 * 4 * 4 * 2 == 0) return false;
  // factorize.h:7-67: divide out the implemented factors, keep the remainder as last factor.
  if (n == 1) return false;  // factors = {1}; 1 is not in the list
  int ntest = n;
  int last = 0;
  for (int i = 0; kFactors[i] && ntest != 1; ++i) {
    int f = kFactors[i];
    while (ntest % f == 0) { ntest /= f; last = f; }
  }
  if (ntest != 1) last = ntest;
  for (int i = 0; kFactors[i]; ++i)
    if (last == kFactors[i]) return true;
  return false;
}

// factorize.h:92-104
int find_closest_factor(int n) {
  if (is_optimal(n)) return n;
  int j = n + 1;
  while (!is_optimal(j)) ++j;
  return j;
}

// Convolution_Mode ordinals, convolution_fftw.h:17-27
enum Mode {
  LINEAR_FULL = 0, LINEAR_SAME_UNPADDED, LINEAR_SAME, LINEAR_VALID,
  CIRCULAR_SAME, CIRCULAR_SAME_PADDED, CIRCULAR_FULL_UNPADDED, CIRCULAR_FULL
};

struct Shape { int h_fft, w_fft, h_dst, w_dst; };

// convolution_fftw.h:52-132
bool plan_shape(int mode, int hs, int ws, int hk, int wk, Shape& s) {
  switch (mode) {
    case LINEAR_FULL:
      s.h_fft = find_closest_factor(hs + hk - 1); s.w_fft = find_closest_factor(ws + wk - 1);
      s.h_dst = hs + hk - 1; s.w_dst = ws + wk - 1; break;
    case LINEAR_SAME_UNPADDED:
      s.h_fft = hs + int(hk / 2.0); s.w_fft = ws + int(wk / 2.0);
      s.h_dst = hs; s.w_dst = ws; break;
    case LINEAR_SAME:
      s.h_fft = find_closest_factor(hs + int(hk / 2.0));
      s.w_fft = find_closest_factor(ws + int(wk / 2.0));
      s.h_dst = hs; s.w_dst = ws; break;
    case LINEAR_VALID:
      if (hk > hs || wk > ws) { s.h_fft = s.w_fft = s.h_dst = s.w_dst = 0; }
      else {
        s.h_fft = find_closest_factor(hs); s.w_fft = find_closest_factor(ws);
        s.h_dst = hs - hk + 1; s.w_dst = ws - wk + 1;
      }
      break;
    case CIRCULAR_SAME:
      s.h_dst = hs; s.w_dst = ws; s.h_fft = hs; s.w_fft = ws; break;
    case CIRCULAR_SAME_PADDED:
      s.h_fft = find_closest_factor(hs + hk); s.w_fft = find_closest_factor(ws + wk);
      s.h_dst = hs; s.w_dst = ws; break;
    case CIRCULAR_FULL_UNPADDED:
      s.h_dst = hs + hk - 1; s.w_dst = ws + wk - 1;
      s.h_fft = find_closest_factor(hs + hk - 1); s.w_fft = find_closest_factor(ws + wk - 1); break;
    case CIRCULAR_FULL:
      s.h_dst = hs + hk - 1; s.w_dst = ws + wk - 1; s.h_fft = s.h_dst; s.w_fft = s.w_dst; break;
    default:
      return false;
  }
  return true;
}

// ---------------------------------------------------------------- FFT path (convolution_fftw.h:169-284)
template <class T>
struct FftwPathPlan {
  Shape s;
  int mode, hs, ws, hk, wk;
  orc::Fft2D<T>* fft = nullptr;
  std::vector<std::complex<T>> a, b;
  FftwPathPlan(int mode_, int hs_, int ws_, int hk_, int wk_)
      : mode(mode_), hs(hs_), ws(ws_), hk(hk_), wk(wk_) {
    plan_shape(mode, hs, ws, hk, wk, s);
    if (s.h_fft > 0 && s.w_fft > 0) {
      fft = new orc::Fft2D<T>(s.h_fft, s.w_fft);
      a.resize((size_t)s.h_fft * s.w_fft);
      b.resize((size_t)s.h_fft * s.w_fft);
    }
  }
  ~FftwPathPlan() { delete fft; }

  void run(const double* src, const double* kernel, double* dst) {
    if (s.h_fft <= 0 || s.w_fft <= 0) return;  // :218-219
    const int H = s.h_fft, W = s.w_fft;
    // :174-177 zero, :181-187 wrap-accumulate
    std::fill(a.begin(), a.end(), std::complex<T>(0, 0));
    std::fill(b.begin(), b.end(), std::complex<T>(0, 0));
    for (int i = 0; i < hs; ++i)
      for (int j = 0; j < ws; ++j) a[(size_t)(i % H) * W + (j % W)] += (T)src[(size_t)i * ws + j];
    for (int i = 0; i < hk; ++i)
      for (int j = 0; j < wk; ++j) b[(size_t)(i % H) * W + (j % W)] += (T)kernel[(size_t)i * wk + j];
    // :190-191 two forward r2c transforms (computed here as full complex DFTs of real data)
    fft->exec(a.data(), -1);
    fft->exec(b.data(), -1);
    // :195-204 product over the h x (w/2+1) half spectrum, written into the kernel's spectrum
    const int Wh = W / 2 + 1;
    for (int i = 0; i < H; ++i)
      for (int j = 0; j < Wh; ++j) b[(size_t)i * W + j] = a[(size_t)i * W + j] * b[(size_t)i * W + j];
    // :208 c2r: the missing half is the Hermitian mirror of the stored half
    for (int i = 0; i < H; ++i)
      for (int j = Wh; j < W; ++j)
        b[(size_t)i * W + j] = std::conj(b[(size_t)((H - i) % H) * W + (W - j)]);
    fft->exec(b.data(), +1);
    // :210-211 scale by 1/(H*W)
    const T scale = (T)1 / (T)((long long)H * W);
    // :227-283 crop
    int oh = 0, ow = 0;
    if (mode == LINEAR_SAME || mode == LINEAR_SAME_UNPADDED) { oh = int(hk / 2.0); ow = int(wk / 2.0); }
    else if (mode == LINEAR_VALID) { oh = hk - 1; ow = wk - 1; }
    for (int i = 0; i < s.h_dst; ++i)
      for (int j = 0; j < s.w_dst; ++j)
        dst[(size_t)i * s.w_dst + j] = (double)(b[(size_t)(i + oh) * W + (j + ow)].real() * scale);
  }
};

// ---------------------------------------------------------------- nested loops (convolution_std.h:86-199)
enum StdMode { S_LINEAR_FULL = 0, S_LINEAR_SAME, S_LINEAR_VALID, S_CIRCULAR_SAME, S_CIRCULAR_FULL };

void std_shape(int mode, int hs, int ws, int hk, int wk, int& hd, int& wd) {
  switch (mode) {  // convolution_std.h:36-58
    case S_LINEAR_FULL: case S_CIRCULAR_FULL: hd = hs + hk - 1; wd = ws + wk - 1; break;
    case S_LINEAR_SAME: case S_CIRCULAR_SAME: hd = hs; wd = ws; break;
    case S_LINEAR_VALID: hd = hs - hk + 1; wd = ws - wk + 1; break;
    default: hd = wd = 0;
  }
}

template <class ACC>
void std_convolve(int mode, int hs, int ws, int hk, int wk, const double* src, const double* kernel,
                  double* dst) {
  int hd, wd;
  std_shape(mode, hs, ws, hk, wk, hd, wd);
  if (hd <= 0 || wd <= 0) return;  // :93-94
  switch (mode) {
    case S_LINEAR_FULL:  // :98-118
      for (int i = 0; i < hd; ++i) {
        int lk = std::max(0, i - hk + 1), hk_ = std::min(hs - 1, i);
        for (int j = 0; j < wd; ++j) {
          int ll = std::max(0, j - wk + 1), hl = std::min(ws - 1, j);
          ACC t = 0;
          for (int k = lk; k <= hk_; ++k)
            for (int l = ll; l <= hl; ++l)
              t += (ACC)src[(size_t)k * ws + l] * (ACC)kernel[(size_t)(i - k) * wk + (j - l)];
          dst[(size_t)i * wd + j] = (double)t;
        }
      }
      break;
    case S_LINEAR_SAME: {  // :119-140
      const int ch = int(hk / 2.0), cw = int(wk / 2.0);
      for (int i = 0; i < hd; ++i) {
        int lk = std::max(0, i - int((hk - 1.0) / 2.0)), hk_ = std::min(hs - 1, i + ch);
        for (int j = 0; j < wd; ++j) {
          int ll = std::max(0, j - int((wk - 1.0) / 2.0)), hl = std::min(ws - 1, j + cw);
          ACC t = 0;
          for (int k = lk; k <= hk_; ++k)
            for (int l = ll; l <= hl; ++l)
              t += (ACC)src[(size_t)k * ws + l] * (ACC)kernel[(size_t)(i - k + ch) * wk + (j - l + cw)];
          dst[(size_t)i * wd + j] = (double)t;
        }
      }
      break;
    }
    case S_LINEAR_VALID:  // :141-159
      for (int i = 0; i < hd; ++i)
        for (int j = 0; j < wd; ++j) {
          ACC t = 0;
          for (int k = i; k <= i + hk - 1; ++k)
            for (int l = j; l <= j + wk - 1; ++l)
              t += (ACC)src[(size_t)k * ws + l] * (ACC)kernel[(size_t)(i + hk - 1 - k) * wk + (j + wk - 1 - l)];
          dst[(size_t)i * wd + j] = (double)t;
        }
      break;
    case S_CIRCULAR_SAME:
    case S_CIRCULAR_FULL:  // :160-188
      for (int i = 0; i < hd; ++i)
        for (int j = 0; j < wd; ++j) {
          ACC t = 0;
          for (int k = 0; k < hk; ++k) {
            int is = (i - k) % hd;
            if (is < 0) is += hd;
            for (int l = 0; l < wk; ++l) {
              int js = (j - l) % wd;
              if (js < 0) js += wd;
              if (is >= 0 && is < hs && js >= 0 && js < ws)
                t += (ACC)src[(size_t)is * ws + js] * (ACC)kernel[(size_t)k * wk + l];
            }
          }
          dst[(size_t)i * wd + j] = (double)t;
        }
      break;
  }
}

// Every FFTW-header mode from nested loops, via the identities of SURVEY.md 8a-8:
//   CIRCULAR_FULL == CIRCULAR_FULL_UNPADDED == LINEAR_FULL; CIRCULAR_SAME_PADDED == top-left
//   hs x ws crop of LINEAR_FULL; LINEAR_SAME_UNPADDED == LINEAR_SAME; CIRCULAR_SAME as is.
template <class ACC>
void direct8(int mode, int hs, int ws, int hk, int wk, const double* src, const double* kernel,
             double* dst) {
  switch (mode) {
    case LINEAR_FULL: case CIRCULAR_FULL: case CIRCULAR_FULL_UNPADDED:
      std_convolve<ACC>(S_LINEAR_FULL, hs, ws, hk, wk, src, kernel, dst); break;
    case LINEAR_SAME: case LINEAR_SAME_UNPADDED:
      std_convolve<ACC>(S_LINEAR_SAME, hs, ws, hk, wk, src, kernel, dst); break;
    case LINEAR_VALID:
      if (hk > hs || wk > ws) return;
      std_convolve<ACC>(S_LINEAR_VALID, hs, ws, hk, wk, src, kernel, dst); break;
    case CIRCULAR_SAME:
      std_convolve<ACC>(S_CIRCULAR_SAME, hs, ws, hk, wk, src, kernel, dst); break;
    case CIRCULAR_SAME_PADDED: {
      const int hf = hs + hk - 1, wf = ws + wk - 1;
      std::vector<double> full((size_t)hf * wf);
      std_convolve<ACC>(S_LINEAR_FULL, hs, ws, hk, wk, src, kernel, full.data());
      for (int i = 0; i < hs; ++i)
        for (int j = 0; j < ws; ++j) dst[(size_t)i * ws + j] = full[(size_t)i * wf + j];
      break;
    }
  }
}

}  // namespace

extern "C" {

int orc_is_optimal(int n) { return is_optimal(n) ? 1 : 0; }
int orc_find_closest_factor(int n) { return find_closest_factor(n); }

int orc_plan(int mode, int hs, int ws, int hk, int wk, int* h_fft, int* w_fft, int* h_dst, int* w_dst) {
  Shape s;
  if (!plan_shape(mode, hs, ws, hk, wk, s)) return 1;
  *h_fft = s.h_fft; *w_fft = s.w_fft; *h_dst = s.h_dst; *w_dst = s.w_dst;
  return 0;
}

// Restatement of init_workspace + convolve of convolution_fftw.h with a long-double DFT.
int orc_convolve_fftw(int mode, int hs, int ws, int hk, int wk, const double* src,
                      const double* kernel, double* dst) {
  if (mode < 0 || mode > 7) return 1;
  FftwPathPlan<long double> p(mode, hs, ws, hk, wk);
  p.run(src, kernel, dst);
  return 0;
}

// Same, in double, over a batch, on `threads` host threads -- the "port" CPU baseline.
// src is [n][hs][ws]; kernel is [hk][wk] (shared) or [n][hk][wk]; dst is [n][h_dst][w_dst].
int orc_convolve_fftw_batch(int mode, int hs, int ws, int hk, int wk, const double* src,
                            const double* kernel, int kernel_is_shared, double* dst, long n,
                            int threads) {
  if (mode < 0 || mode > 7) return 1;
  Shape s;
  plan_shape(mode, hs, ws, hk, wk, s);
  if (threads < 1) threads = 1;
  auto work = [&](int t) {
    FftwPathPlan<double> p(mode, hs, ws, hk, wk);  // one Workspace per thread, like the reference
    for (long i = t; i < n; i += threads)
      p.run(src + (size_t)i * hs * ws, kernel + (kernel_is_shared ? 0 : (size_t)i * hk * wk),
            dst + (size_t)i * s.h_dst * s.w_dst);
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < threads; ++t) pool.emplace_back(work, t);
  work(0);
  for (auto& th : pool) th.join();
  return 0;
}

int orc_std_shape(int std_mode, int hs, int ws, int hk, int wk, int* hd, int* wd) {
  if (std_mode < 0 || std_mode > 4) return 1;
  std_shape(std_mode, hs, ws, hk, wk, *hd, *wd);
  return 0;
}

// convolution_std.h restated with the reference's own double accumulation ...
int orc_convolve_std(int std_mode, int hs, int ws, int hk, int wk, const double* src,
                     const double* kernel, double* dst) {
  if (std_mode < 0 || std_mode > 4) return 1;
  std_convolve<double>(std_mode, hs, ws, hk, wk, src, kernel, dst);
  return 0;
}

// ... and with long-double accumulation (tighter ground truth for the 1e-12 bar).
int orc_convolve_std_ld(int std_mode, int hs, int ws, int hk, int wk, const double* src,
                        const double* kernel, double* dst) {
  if (std_mode < 0 || std_mode > 4) return 1;
  std_convolve<long double>(std_mode, hs, ws, hk, wk, src, kernel, dst);
  return 0;
}

// All eight convolution_fftw.h modes from nested loops (long-double accumulation).
int orc_convolve_direct8(int mode, int hs, int ws, int hk, int wk, const double* src,
                         const double* kernel, double* dst) {
  if (mode < 0 || mode > 7) return 1;
  direct8<long double>(mode, hs, ws, hk, wk, src, kernel, dst);
  return 0;
}

// Batched nested loops (double, like the reference) on `threads` host threads.
int orc_convolve_direct8_batch(int mode, int hs, int ws, int hk, int wk, const double* src,
                               const double* kernel, int kernel_is_shared, double* dst, long n,
                               int threads) {
  if (mode < 0 || mode > 7) return 1;
  Shape s;
  plan_shape(mode, hs, ws, hk, wk, s);
  if (threads < 1) threads = 1;
  auto work = [&](int t) {
    for (long i = t; i < n; i += threads)
      direct8<double>(mode, hs, ws, hk, wk, src + (size_t)i * hs * ws,
                      kernel + (kernel_is_shared ? 0 : (size_t)i * hk * wk),
                      dst + (size_t)i * s.h_dst * s.w_dst);
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < threads; ++t) pool.emplace_back(work, t);
  work(0);
  for (auto& th : pool) th.join();
  return 0;
}

}  // extern "C"
