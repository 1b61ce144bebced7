// TEST INFRASTRUCTURE ONLY -- implementation of the nine-name fftw3.h stand-in (see fftw3.h here).
#include "fftw3.h"

#include <complex>
#include <cstdlib>
#include <vector>

#include "../mixed_fft.h"

struct orc_fftw_plan_s {
  int h, w, wh;
  bool r2c;
  double* real;
  std::complex<double>* cplx;
  orc::Fft1D<double> fh, fw;
  std::vector<std::complex<double>> a, b;
};

extern "C" {

void* fftw_malloc(size_t n) { return std::malloc(n ? n : 1); }
void fftw_free(void* p) { std::free(p); }

static fftw_plan make_plan(int n0, int n1, double* re, fftw_complex* cx, bool r2c) {
  orc_fftw_plan_s* p = new orc_fftw_plan_s();
  p->h = n0; p->w = n1; p->wh = n1 / 2 + 1; p->r2c = r2c;
  p->real = re; p->cplx = reinterpret_cast<std::complex<double>*>(cx);
  if (n0 > 0 && n1 > 0) {
    p->fh.init(n0); p->fw.init(n1);
    p->a.resize(n0 > n1 ? n0 : n1); p->b.resize(p->a.size());
  }
  return p;
}

fftw_plan fftw_plan_dft_r2c_2d(int n0, int n1, double* in, fftw_complex* out, unsigned) {
  return make_plan(n0, n1, in, out, true);
}
fftw_plan fftw_plan_dft_c2r_2d(int n0, int n1, fftw_complex* in, double* out, unsigned) {
  return make_plan(n0, n1, out, in, false);
}

void fftw_execute(const fftw_plan p) {
  const int h = p->h, w = p->w, wh = p->wh;
  if (h <= 0 || w <= 0) return;
  typedef std::complex<double> C;
  C* a = p->a.data();
  C* b = p->b.data();
  if (p->r2c) {
    // rows: two real rows ride one complex transform (z = x_i + i*x_{i+1}), separated afterwards by
    // Hermitian symmetry: A[j] = (Z[j] + conj Z[w-j]) / 2, B[j] = (Z[j] - conj Z[w-j]) / (2i)
    for (int i = 0; i < h; i += 2) {
      const bool two = i + 1 < h;
      const double* r0 = p->real + (size_t)i * w;
      const double* r1 = r0 + w;
      for (int j = 0; j < w; ++j) a[j] = C(r0[j], two ? r1[j] : 0.0);
      p->fw.exec(a, b, -1);
      C* o0 = p->cplx + (size_t)i * wh;
      C* o1 = o0 + wh;
      for (int j = 0; j < wh; ++j) {
        const C z = a[j], zc = std::conj(a[j ? w - j : 0]);
        o0[j] = 0.5 * (z + zc);
        if (two) { const C d = z - zc; o1[j] = C(0.5 * d.imag(), -0.5 * d.real()); }
      }
    }
    for (int j = 0; j < wh; ++j) {  // columns
      for (int i = 0; i < h; ++i) a[i] = p->cplx[(size_t)i * wh + j];
      p->fh.exec(a, b, -1);
      for (int i = 0; i < h; ++i) p->cplx[(size_t)i * wh + j] = a[i];
    }
  } else {  // c2r, destroys its input like FFTW does
    for (int j = 0; j < wh; ++j) {
      for (int i = 0; i < h; ++i) a[i] = p->cplx[(size_t)i * wh + j];
      p->fh.exec(a, b, +1);
      for (int i = 0; i < h; ++i) p->cplx[(size_t)i * wh + j] = a[i];
    }
    // rows: two Hermitian half-spectra ride one complex backward transform (W = Y0 + i*Y1)
    for (int i = 0; i < h; i += 2) {
      const bool two = i + 1 < h;
      const C* y0 = p->cplx + (size_t)i * wh;
      const C* y1 = y0 + wh;
      for (int j = 0; j < w; ++j) {
        const C u = j < wh ? y0[j] : std::conj(y0[w - j]);
        const C v = two ? (j < wh ? y1[j] : std::conj(y1[w - j])) : C(0, 0);
        a[j] = C(u.real() - v.imag(), u.imag() + v.real());
      }
      p->fw.exec(a, b, +1);
      double* r0 = p->real + (size_t)i * w;
      double* r1 = r0 + w;
      for (int j = 0; j < w; ++j) { r0[j] = a[j].real(); if (two) r1[j] = a[j].imag(); }
    }
  }
}

void fftw_destroy_plan(fftw_plan p) { delete p; }

}  // extern "C"
