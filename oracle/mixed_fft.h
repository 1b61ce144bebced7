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
    typedef std::complex<T> C;
    if (n <= 1) return;
    C* x = a;
    C* y = b;
    int len = n, s = 1;
    C t[64], wp[64];
    std::vector<C> wr_big;
    for (size_t st = 0; st < radices.size(); ++st) {
      const int r = radices[st];
      const int m = len / r;
      const int wr_step = n / r;    // w_r^1 = tw[n/r]
      const int wn_step = n / len;  // w_len^1 = tw[n/len]
      C wr_small[13 * 13];
      C* wr = wr_small;
      C* tt = t;
      C* wpp = wp;
      std::vector<C> big_t, big_wp;
      if (r > 13) {
        wr_big.resize((size_t)r * r); wr = wr_big.data();
        if (r > 64) { big_t.resize(r); big_wp.resize(r); tt = big_t.data(); wpp = big_wp.data(); }
      }
      if (r > 5 || r == 1)
        for (int k = 0; k < r; ++k)
          for (int i = 0; i < r; ++i) {
            C w = tw[(size_t)((long long)i * k % r) * wr_step];
            wr[(size_t)k * r + i] = sign > 0 ? std::conj(w) : w;
          }
      const T sg = (T)(sign > 0 ? 1 : -1);   // multiply by sg*i rotates the way this direction needs
      const T s3 = (T)0.866025403784438646763723170752936183L;
      const T c51 = (T)0.309016994374947424102293417182819059L, c52 = (T)-0.809016994374947424102293417182819059L;
      const T s51 = (T)0.951056516295153572116439333379382143L, s52 = (T)0.587785252292473129168705954639072769L;
      for (int p = 0; p < m; ++p) {
        for (int k = 1; k < r; ++k) {
          C w = tw[(size_t)p * k * wn_step];
          wpp[k] = sign > 0 ? std::conj(w) : w;
        }
        for (int q = 0; q < s; ++q) {
          const C* xi = x + q + s * p;
          C* yo = y + q + s * r * p;
          const int xs = s * m;
          if (r == 2) {
            const C u = xi[0], v = xi[xs];
            yo[0] = u + v;
            yo[s] = (u - v) * wpp[1];
          } else if (r == 3) {
            const C u = xi[0], v = xi[xs], w = xi[2 * xs];
            const C tsum = v + w, d = v - w;
            const C mm = u - (T)0.5 * tsum;
            const C rot(-sg * s3 * d.imag(), sg * s3 * d.real());   // sg * i * s3 * d
            yo[0] = u + tsum;
            yo[s] = (mm + rot) * wpp[1];
            yo[2 * s] = (mm - rot) * wpp[2];
          } else if (r == 4) {
            const C u0 = xi[0], u1 = xi[xs], u2 = xi[2 * xs], u3 = xi[3 * xs];
            const C a0 = u0 + u2, a1 = u0 - u2, a2 = u1 + u3, d = u1 - u3;
            const C rot(-sg * d.imag(), sg * d.real());             // sg * i * d
            yo[0] = a0 + a2;
            yo[s] = (a1 + rot) * wpp[1];
            yo[2 * s] = (a0 - a2) * wpp[2];
            yo[3 * s] = (a1 - rot) * wpp[3];
          } else if (r == 5) {
            const C u0 = xi[0], u1 = xi[xs], u2 = xi[2 * xs], u3 = xi[3 * xs], u4 = xi[4 * xs];
            const C t1 = u1 + u4, t2 = u2 + u3, t3 = u1 - u4, t4 = u2 - u3;
            const C m1 = u0 + c51 * t1 + c52 * t2, m2 = u0 + c52 * t1 + c51 * t2;
            const C n1 = s51 * t3 + s52 * t4, n2 = s52 * t3 - s51 * t4;
            const C r1(-sg * n1.imag(), sg * n1.real()), r2(-sg * n2.imag(), sg * n2.real());   // sg * i * n
            yo[0] = u0 + t1 + t2;
            yo[s] = (m1 + r1) * wpp[1];
            yo[4 * s] = (m1 - r1) * wpp[4];
            yo[2 * s] = (m2 + r2) * wpp[2];
            yo[3 * s] = (m2 - r2) * wpp[3];
          } else {
            for (int i = 0; i < r; ++i) tt[i] = xi[i * xs];
            for (int k = 0; k < r; ++k) {
              const C* wk = &wr[(size_t)k * r];
              C acc = tt[0];
              for (int i = 1; i < r; ++i) acc += tt[i] * wk[i];
              yo[k * s] = k ? acc * wpp[k] : acc;
            }
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
