// Development aid: steps the five passes of the period-108 kernel (conv_pfa108.cuh) thread by thread on the CPU
// and compares with a long-double direct convolution, so that index maps can be debugged without a GPU.
//   nvcc -O1 -std=c++17 -I fftconvolution_b200/csrc tools/emulate_pfa108.cu -o /tmp/emu && /tmp/emu
// Not part of the library; nothing here is a product path.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "conv_pfa108.cuh"

using namespace b2c;
using namespace b2c::pfa;

static std::vector<cplx> host_kspec(const std::vector<double>& ker, int hk, int wk) {
  std::vector<cplx> out((size_t)P * KK);
  const long double two_pi = 6.283185307179586476925286766559005768L;
  for (int pos = 0; pos < P; ++pos)
    for (int kk = 0; kk < KK; ++kk) {
      const int kv = kv_of(kk), kh = kh_of(pos / 27, pos % 27);
      long double re = 0, im = 0;
      for (int a = 0; a < hk; ++a)
        for (int b = 0; b < wk; ++b) {
          const long double ang = -two_pi * (long double)((kv * (a % P) + kh * (b % P)) % P) / P;
          re += ker[a * wk + b] * cosl(ang);
          im += ker[a * wk + b] * sinl(ang);
        }
      const long double sc = 0.5L / (108.0L * 108.0L);
      out[(size_t)pos * KK + kk] = make_double2((double)(re * sc), (double)(im * sc));
    }
  return out;
}

static double run_case(int hs, int ws, int hk, int wk, int off_h, int off_w, int hd, int wd, unsigned seed, bool circ = false) {
  srand(seed);
  std::vector<double> src((size_t)hs * ws), ker((size_t)hk * wk), dst((size_t)hd * wd, NAN);
  for (auto& v : src) v = rand() / (double)RAND_MAX;
  for (auto& v : ker) v = rand() / (double)RAND_MAX;
  std::vector<cplx> ks = host_kspec(ker, hk, wk);
  std::vector<cplx> Z((size_t)HALF * LD, make_double2(NAN, NAN));
  Params p;
  p.hs = hs; p.ws = ws; p.hd = hd; p.wd = wd; p.off_h = off_h; p.off_w = off_w;
  p.hs_in = hs; p.ws_in = ws; p.sh_h = 0; p.sh_w = 0;
  if (circ) {   // CIRCULAR_SAME as the linear convolution of the periodically extended source, cropped at n_ker - 1
    p.sh_h = hk - 1; p.sh_w = wk - 1; p.hs_in = hs + hk - 1; p.ws_in = ws + wk - 1;
  }
  p.src = src.data(); p.dst = dst.data(); p.n_images = 1; p.kspec = ks.data();
  cplx v[27];
  for (int t = 0; t < THREADS; ++t) {
    const VThread q = vthread(t);
    if (q.act) { if (circ) p1_load<true>(v, q, p, src.data()); else p1_load<false>(v, q, p, src.data()); dft27(v); p1_store(v, q, Z.data()); }
  }
  for (int t = 0; t < THREADS; ++t) pass2(t, Z.data());
  for (int t = 0; t < THREADS; ++t) {
    cplx* z3 = p3_base(t, Z.data());
    p3_load(v, z3); dft27(v); p3_mul(v, t, ks.data()); dft27(v); p3_store(v, z3);
  }
  for (int t = 0; t < THREADS; ++t) pass4(t, Z.data());
  for (int t = 0; t < THREADS; ++t) {
    const VThread q = vthread(t);
    if (q.act) { p5_load(v, q, Z.data()); dft27(v); p5_store(v, q, p, dst.data()); }
  }
  long double num = 0, den = 0;
  for (int i = 0; i < hd; ++i)
    for (int j = 0; j < wd; ++j) {
      long double acc = 0;
      for (int a = 0; a < hk; ++a)
        for (int b = 0; b < wk; ++b) {
          int r = i + off_h - a, c = j + off_w - b;
          if (circ) { r = ((i - a) % hs + hs) % hs; c = ((j - b) % ws + ws) % ws; }
          if (r >= 0 && r < hs && c >= 0 && c < ws) acc += (long double)ker[a * wk + b] * src[r * ws + c];
        }
      const long double d = dst[i * wd + j] - acc;
      num += d * d; den += acc * acc;
    }
  return (double)sqrtl(num / den);
}

int main() {
  struct { int hs, ws, hk, wk, off_h, off_w, hd, wd; } cases[] = {
      {100, 100, 15, 15, 7, 7, 100, 100},   // LINEAR_SAME headline
      {94, 94, 15, 15, 0, 0, 108, 108},     // LINEAR_FULL at exactly the period
      {60, 101, 9, 8, 0, 0, 68, 108},       // rectangle
      {108, 108, 5, 3, 4, 2, 104, 106},     // LINEAR_VALID
      {1, 50, 1, 7, 0, 3, 1, 50},           // one row
      {50, 1, 7, 1, 3, 0, 50, 1},           // one column
  };
  int bad = 0;
  {
    const double e = run_case(97, 89, 9, 3, 8, 2, 97, 89, 77, true);
    printf(" 97x 89 (*)  9x 3 CIRCULAR_SAME by periodic extension : rel err %.3e %s\n", e, e <= 1e-13 ? "ok" : "FAIL");
    bad += !(e <= 1e-13);
  }
  for (auto& c : cases) {
    const double e = run_case(c.hs, c.ws, c.hk, c.wk, c.off_h, c.off_w, c.hd, c.wd, 1234);
    printf("%3dx%3d (*) %2dx%2d off (%d,%d) -> %3dx%3d : rel err %.3e %s\n", c.hs, c.ws, c.hk, c.wk, c.off_h, c.off_w,
           c.hd, c.wd, e, e <= 1e-13 ? "ok" : "FAIL");
    bad += !(e <= 1e-13);
  }
  return bad;
}
