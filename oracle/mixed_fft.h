// TEST INFRASTRUCTURE ONLY -- part of oracle/, never linked into the product library.
//
// A small any-length complex FFT used (a) by the oracle restatement of
// FFTW_Convolution::fftw_circular_convolution (reference Convolution/src/convolution_fftw.h:169-214)
// where the reference calls libfftw3, and (b) by the stand-in fftw3.h shim under oracle/shim/
// that lets the UNMODIFIED reference header compile in a container without libfftw3.
//
// Third-party dependency being stood in for: FFTW3 (system library, no pinned version in the
// reference tree; the write-up used 3.2.2, Convolution/Doc/latex.tex:68).  Published semantics
// restated here: fftw_plan_dft_r2c_2d / c2r_2d compute the UNNORMALISED forward (sign -1) and
// backward (sign +1) DFT, r2c output is h x (w/2+1) interleaved complex, last axis halved.
//
// Algorithm: recursive-free Stockham autosort, decimation in frequency, generic radix-p
// butterflies (O(p^2) each) over the prime factors of n.  Prime lengths degrade to the naive
// DFT, which is what an oracle wants.  Templated on the real type: long double is the
// ground-truth instantiation, double the timed one.
#pragma once
#include <algorithm>
#include <cmath>
#include <complex>
#include <vector>

namespace orc {

template <class T>
struct Fft1D {
  int n = 0;
  std::vector<int> radices;
  std::vector<std::complex<T>> tw;  // tw[k] = exp(-2*pi*i*k/n)

  explicit Fft1D(int n_ = 1) { init(n_); }

  void init(int n_) {
    n = n_;
    radices.clear();
    int rem = n;
    // Small radices first keeps the O(p^2) butterflies cheap; 4 before 2 halves the passes.
    while (rem % 4 == 0) { radices.push_back(4); rem /= 4; }
    for (int p = 2; p * p <= rem; ++p)
      while (rem % p == 0) { radices.push_back(p); rem /= p; }
    if (rem > 1) radices.push_back(rem);
    tw.resize(n > 0 ? n : 1);
    const long double two_pi = 6.283185307179586476925286766559005768L;
    for (int k = 0; k < n; ++k) {
      long double a = two_pi * (long double)k / (long double)n;
      tw[k] = std::complex<T>((T)cosl(a), (T)(-sinl(a)));
    }
  }

  // out-of-place; `a` holds the input and is clobbered; the result ends in `a`.
  // sign = -1 forward, +1 backward (unnormalised both ways, like FFTW).
  void exec(std::complex<T>* a, std::complex<T>* b, int sign) const {
    if (n <= 1) return;
    std::complex<T>* x = a;
    std::complex<T>* y = b;
    int len = n, s = 1;
    std::vector<std::complex<T>> t(16), wr, wp(16);
    for (size_t st = 0; st < radices.size(); ++st) {
      const int r = radices[st];
      const int m = len / r;
      if ((int)t.size() < r) { t.resize(r); wp.resize(r); }
      const int wr_step = n / r;    // w_r^1 = tw[n/r]
      const int wn_step = n / len;  // w_len^1 = tw[n/len]
      wr.resize((size_t)r * r);
      for (int k = 0; k < r; ++k)
        for (int i = 0; i < r; ++i) {
          std::complex<T> w = tw[(size_t)((long long)i * k % r) * wr_step];
          wr[(size_t)k * r + i] = sign > 0 ? std::conj(w) : w;
        }
      for (int p = 0; p < m; ++p) {
        for (int k = 0; k < r; ++k) {
          std::complex<T> w = tw[(size_t)p * k * wn_step];
          wp[k] = sign > 0 ? std::conj(w) : w;
        }
        for (int q = 0; q < s; ++q) {
          for (int i = 0; i < r; ++i) t[i] = x[q + s * (p + m * i)];
          if (r == 2) {
            y[q + s * (2 * p)] = t[0] + t[1];
            y[q + s * (2 * p + 1)] = (t[0] - t[1]) * wp[1];
            continue;
          }
          for (int k = 0; k < r; ++k) {
            const std::complex<T>* wk = &wr[(size_t)k * r];
            std::complex<T> acc = t[0];
            for (int i = 1; i < r; ++i) acc += t[i] * wk[i];
            y[q + s * (r * p + k)] = k ? acc * wp[k] : acc;
          }
        }
      }
      std::swap(x, y);
      len = m;
      s *= r;
    }
    if (x != a)
      for (int i = 0; i < n; ++i) a[i] = x[i];
  }
};

// Unnormalised 2-D complex DFT, row-major h x w, in place.
template <class T>
struct Fft2D {
  int h, w;
  Fft1D<T> fh, fw;
  Fft2D(int h_, int w_) : h(h_), w(w_), fh(h_), fw(w_) {}
  void exec(std::complex<T>* d, int sign) const {
    std::vector<std::complex<T>> a(std::max(h, w)), b(std::max(h, w));
    for (int i = 0; i < h; ++i) {
      for (int j = 0; j < w; ++j) a[j] = d[(size_t)i * w + j];
      fw.exec(a.data(), b.data(), sign);
      for (int j = 0; j < w; ++j) d[(size_t)i * w + j] = a[j];
    }
    for (int j = 0; j < w; ++j) {
      for (int i = 0; i < h; ++i) a[i] = d[(size_t)i * w + j];
      fh.exec(a.data(), b.data(), sign);
      for (int i = 0; i < h; ++i) d[(size_t)i * w + j] = a[i];
    }
  }
};

}  // namespace orc
