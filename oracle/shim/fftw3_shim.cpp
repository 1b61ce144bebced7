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
  std::complex<double>* a = p->a.data();
  std::complex<double>* b = p->b.data();
  if (p->r2c) {
    for (int i = 0; i < h; ++i) {  // rows: real -> half spectrum
      for (int j = 0; j < w; ++j) a[j] = p->real[(size_t)i * w + j];
      p->fw.exec(a, b, -1);
      for (int j = 0; j < wh; ++j) p->cplx[(size_t)i * wh + j] = a[j];
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
    for (int i = 0; i < h; ++i) {
      for (int j = 0; j < wh; ++j) a[j] = p->cplx[(size_t)i * wh + j];
      for (int j = wh; j < w; ++j) a[j] = std::conj(p->cplx[(size_t)i * wh + (w - j)]);
      p->fw.exec(a, b, +1);
      for (int j = 0; j < w; ++j) p->real[(size_t)i * w + j] = a[j].real();
    }
  }
}

void fftw_destroy_plan(fftw_plan p) { delete p; }

}  // extern "C"
