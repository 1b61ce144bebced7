// Device-side FFT building blocks of the generic (any 2/3/5/7-smooth size) convolution path.
//
// Transform scheme: IN-PLACE decimation in frequency for the forward transform (natural order
// in, digit-scrambled order out) and its exact mirror, decimation in time, for the backward
// transform (scrambled in, natural out).  A convolution multiplies two spectra that were
// scrambled by the same schedule, so no reordering pass is ever needed and one work buffer per
// image is enough (no Stockham ping-pong copy) -- half the shared memory of an autosort FFT.
//
// Replaces the three fftw_execute calls of the reference (convolution_fftw.h:190-191,208).
#pragma once
#include <cuda_runtime.h>

namespace b2c {

typedef double2 cplx;

// complex arithmetic and register butterflies are plain math: also compilable for the host (tools/emulate_pfa108.cu
// steps the headline kernel's passes thread by thread on the CPU as a development aid -- never a product path)
#define B2C_MATH __host__ __device__ __forceinline__

B2C_MATH cplx cadd(cplx a, cplx b) { return make_double2(a.x + b.x, a.y + b.y); }
B2C_MATH cplx csub(cplx a, cplx b) { return make_double2(a.x - b.x, a.y - b.y); }
B2C_MATH cplx cmul(cplx a, cplx b) {
  return make_double2(fma(a.x, b.x, -a.y * b.y), fma(a.x, b.y, a.y * b.x));
}
// a * conj(b)
B2C_MATH cplx cmulc(cplx a, cplx b) {
  return make_double2(fma(a.x, b.x, a.y * b.y), fma(a.y, b.x, -a.x * b.y));
}
// -i * a
B2C_MATH cplx mul_mi(cplx a) { return make_double2(a.y, -a.x); }
// +i * a
B2C_MATH cplx mul_pi(cplx a) { return make_double2(-a.y, a.x); }
B2C_MATH cplx cswap(cplx a) { return make_double2(a.y, a.x); }

// One thread asks the copy engine to bring `bytes` of global memory into L2 (no registers, no waiting): the
// next image of a persistent CTA, so that its first pass meets L2 latency instead of HBM latency.
__device__ __forceinline__ void l2_prefetch(const void* p, size_t bytes) {
  const unsigned n = (unsigned)(bytes & ~(size_t)15);
  if ((reinterpret_cast<size_t>(p) & 15) == 0 && n >= 16)   // bulk operations need 16-byte alignment
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(n) : "memory");
}

// ---- magnitude balance of packed transforms ------------------------------------------------------
// Kernels that transform a source and a kernel as ONE complex image (z = s + i*k, then a Hermitian split)
// extract the kernel's spectrum as a difference of source-sized numbers: relative accuracy degrades by
// max|s| / max|k|.  The reference transforms the two separately (convolution_fftw.h:190-191).  Cure: pack
// z = s + i*alpha*k with alpha the power of two that equalises the magnitudes (exact), undo it in the scale.
__device__ __forceinline__ double balance_from_maxima(double ms, double mk) {
  if (!(ms > 0.0) || !(mk > 0.0) || !isfinite(ms) || !isfinite(mk)) return 1.0;
  return exp2((double)(ilogb(ms) - ilogb(mk)));
}

// The whole CTA scans one image and its kernel (the data is read again right after: L2 hits) and returns alpha
// to every thread.  `scratch`: 64 doubles any thread of the CTA can write (the still unused work buffer).
// Contains two __syncthreads().
__device__ __forceinline__ double cta_balance(const double* __restrict__ a, int na, const double* __restrict__ k, int nk,
                                              double* scratch) {
  double ms = 0.0, mk = 0.0;
  constexpr int U = 8;   // independent loads in flight per thread: the scan is latency-bound otherwise
  const int step = blockDim.x;
  for (int base = threadIdx.x; base < na; base += U * step) {
    double v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = base + u * step < na ? __ldg(a + base + u * step) : 0.0;
#pragma unroll
    for (int u = 0; u < U; ++u) ms = fmax(ms, fabs(v[u]));
  }
  for (int base = threadIdx.x; base < nk; base += U * step) {
    double v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = base + u * step < nk ? __ldg(k + base + u * step) : 0.0;
#pragma unroll
    for (int u = 0; u < U; ++u) mk = fmax(mk, fabs(v[u]));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    ms = fmax(ms, __shfl_xor_sync(0xffffffffu, ms, o));
    mk = fmax(mk, __shfl_xor_sync(0xffffffffu, mk, o));
  }
  const int w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  if ((threadIdx.x & 31) == 0) { scratch[w] = ms; scratch[32 + w] = mk; }
  __syncthreads();
  ms = 0.0; mk = 0.0;
  for (int i = 0; i < nw; ++i) { ms = fmax(ms, scratch[i]); mk = fmax(mk, scratch[32 + i]); }
  __syncthreads();   // scratch may be overwritten after this
  return balance_from_maxima(ms, mk);
}

// ---- forward DFT butterflies, w = exp(-2*pi*i/R), in registers --------------------------------
template <int R>
struct Bfly;

template <>
struct Bfly<2> {
  static B2C_MATH void run(cplx* v) {
    cplx a = v[0], b = v[1];
    v[0] = cadd(a, b);
    v[1] = csub(a, b);
  }
};

template <>
struct Bfly<3> {
  static B2C_MATH void run(cplx* v) {
    const double s = 0.86602540378443864676;  // sin(2*pi/3)
    cplx t = cadd(v[1], v[2]);
    cplx d = csub(v[1], v[2]);
    cplx m = make_double2(fma(-0.5, t.x, v[0].x), fma(-0.5, t.y, v[0].y));
    v[0] = cadd(v[0], t);
    // X1 = m - i*s*d ; X2 = m + i*s*d
    v[1] = make_double2(fma(s, d.y, m.x), fma(-s, d.x, m.y));
    v[2] = make_double2(fma(-s, d.y, m.x), fma(s, d.x, m.y));
  }
};

template <>
struct Bfly<4> {
  static B2C_MATH void run(cplx* v) {
    cplx a = cadd(v[0], v[2]), b = csub(v[0], v[2]);
    cplx c = cadd(v[1], v[3]), d = mul_mi(csub(v[1], v[3]));
    v[0] = cadd(a, c);
    v[1] = cadd(b, d);
    v[2] = csub(a, c);
    v[3] = csub(b, d);
  }
};

template <>
struct Bfly<5> {
  static B2C_MATH void run(cplx* v) {
    const double c1 = 0.30901699437494742410;   // cos(2*pi/5)
    const double c2 = -0.80901699437494742410;  // cos(4*pi/5)
    const double s1 = 0.95105651629515357212;   // sin(2*pi/5)
    const double s2 = 0.58778525229247312917;   // sin(4*pi/5)
    cplx t1 = cadd(v[1], v[4]), t2 = cadd(v[2], v[3]);
    cplx t3 = csub(v[1], v[4]), t4 = csub(v[2], v[3]);
    cplx x0 = v[0];
    cplx m1 = make_double2(fma(c2, t2.x, fma(c1, t1.x, x0.x)), fma(c2, t2.y, fma(c1, t1.y, x0.y)));
    cplx m2 = make_double2(fma(c1, t2.x, fma(c2, t1.x, x0.x)), fma(c1, t2.y, fma(c2, t1.y, x0.y)));
    cplx n1 = make_double2(fma(s2, t4.x, s1 * t3.x), fma(s2, t4.y, s1 * t3.y));
    cplx n2 = make_double2(fma(-s1, t4.x, s2 * t3.x), fma(-s1, t4.y, s2 * t3.y));
    v[0] = cadd(x0, cadd(t1, t2));
    v[1] = make_double2(m1.x + n1.y, m1.y - n1.x);  // m1 - i n1
    v[4] = make_double2(m1.x - n1.y, m1.y + n1.x);  // m1 + i n1
    v[2] = make_double2(m2.x + n2.y, m2.y - n2.x);
    v[3] = make_double2(m2.x - n2.y, m2.y + n2.x);
  }
};

template <>
struct Bfly<7> {
  static B2C_MATH void run(cplx* v) {
    const double c1 = 0.62348980185873353053;   // cos(2*pi/7)
    const double c2 = -0.22252093395631440429;  // cos(4*pi/7)
    const double c3 = -0.90096886790241912624;  // cos(6*pi/7)
    const double s1 = 0.78183148246802980871;   // sin(2*pi/7)
    const double s2 = 0.97492791218182360702;   // sin(4*pi/7)
    const double s3 = 0.43388373911755812048;   // sin(6*pi/7)
    cplx ta = cadd(v[1], v[6]), tb = cadd(v[2], v[5]), tc = cadd(v[3], v[4]);
    cplx da = csub(v[1], v[6]), db = csub(v[2], v[5]), dc = csub(v[3], v[4]);
    cplx x0 = v[0];
    // symmetric parts: M_m = x0 + c(m)ta + c(2m)tb + c(3m)tc  (indices mod 7, cos even)
    cplx M1 = make_double2(fma(c3, tc.x, fma(c2, tb.x, fma(c1, ta.x, x0.x))),
                           fma(c3, tc.y, fma(c2, tb.y, fma(c1, ta.y, x0.y))));
    cplx M2 = make_double2(fma(c1, tc.x, fma(c3, tb.x, fma(c2, ta.x, x0.x))),
                           fma(c1, tc.y, fma(c3, tb.y, fma(c2, ta.y, x0.y))));
    cplx M3 = make_double2(fma(c2, tc.x, fma(c1, tb.x, fma(c3, ta.x, x0.x))),
                           fma(c2, tc.y, fma(c1, tb.y, fma(c3, ta.y, x0.y))));
    // antisymmetric parts: N_m = s(m)da + s(2m)db + s(3m)dc  (sin odd: s(4)=-s3, s(6)=-s1, s(9)=s2)
    cplx N1 = make_double2(fma(s3, dc.x, fma(s2, db.x, s1 * da.x)),
                           fma(s3, dc.y, fma(s2, db.y, s1 * da.y)));
    cplx N2 = make_double2(fma(-s1, dc.x, fma(-s3, db.x, s2 * da.x)),
                           fma(-s1, dc.y, fma(-s3, db.y, s2 * da.y)));
    cplx N3 = make_double2(fma(s2, dc.x, fma(-s1, db.x, s3 * da.x)),
                           fma(s2, dc.y, fma(-s1, db.y, s3 * da.y)));
    v[0] = cadd(cadd(x0, ta), cadd(tb, tc));
    v[1] = make_double2(M1.x + N1.y, M1.y - N1.x);
    v[6] = make_double2(M1.x - N1.y, M1.y + N1.x);
    v[2] = make_double2(M2.x + N2.y, M2.y - N2.x);
    v[5] = make_double2(M2.x - N2.y, M2.y + N2.x);
    v[3] = make_double2(M3.x + N3.y, M3.y - N3.x);
    v[4] = make_double2(M3.x - N3.y, M3.y + N3.x);
  }
};

template <>
struct Bfly<8> {
  static B2C_MATH void run(cplx* v) {
    const double h = 0.70710678118654752440;  // 1/sqrt(2)
    cplx e[4] = {v[0], v[2], v[4], v[6]};
    cplx o[4] = {v[1], v[3], v[5], v[7]};
    Bfly<4>::run(e);
    Bfly<4>::run(o);
    // o[k] *= w8^k
    cplx o1 = make_double2(h * (o[1].x + o[1].y), h * (o[1].y - o[1].x));   // (1-i)/sqrt2
    cplx o2 = mul_mi(o[2]);
    cplx o3 = make_double2(h * (o[3].y - o[3].x), -h * (o[3].x + o[3].y));  // (-1-i)/sqrt2
    v[0] = cadd(e[0], o[0]); v[4] = csub(e[0], o[0]);
    v[1] = cadd(e[1], o1);   v[5] = csub(e[1], o1);
    v[2] = cadd(e[2], o2);   v[6] = csub(e[2], o2);
    v[3] = cadd(e[3], o3);   v[7] = csub(e[3], o3);
  }
};

// Forward DFT of R points in registers for every radix the tuned kernels use: the plain butterflies above,
// prime-factor compositions of them (6, 10, 12, 14, 18, 20) and Cooley-Tukey 3x3 / 4x4 (9, 16).
template <int R>
struct BigBfly { static B2C_MATH void run(cplx* v) { Bfly<R>::run(v); } };

// ---- small-radix butterflies by prime-factor (Good-Thomas) composition: no internal twiddles ----
__host__ __device__ constexpr int inv_mod(int a, int m) {
  for (int x = 1; x < m; ++x)
    if ((a * x) % m == 1) return x;
  return 1;
}

// N = A * B, gcd(A, B) = 1: input n = (B*n1 + A*n2) mod N, output k = (B*B'*k1 + A*A'*k2) mod N
template <int A, int B>
B2C_MATH void bfly_pfa(cplx* v) {
  constexpr int N = A * B;
  constexpr int ea = B * inv_mod(B % A, A), eb = A * inv_mod(A % B, B);
  cplx u[A][B];
#pragma unroll
  for (int n1 = 0; n1 < A; ++n1) {
    cplx t[B];
#pragma unroll
    for (int n2 = 0; n2 < B; ++n2) t[n2] = v[(B * n1 + A * n2) % N];
    BigBfly<B>::run(t);
#pragma unroll
    for (int k2 = 0; k2 < B; ++k2) u[n1][k2] = t[k2];
  }
#pragma unroll
  for (int k2 = 0; k2 < B; ++k2) {
    cplx s[A];
#pragma unroll
    for (int n1 = 0; n1 < A; ++n1) s[n1] = u[n1][k2];
    BigBfly<A>::run(s);
#pragma unroll
    for (int k1 = 0; k1 < A; ++k1) v[(ea * k1 + eb * k2) % N] = s[k1];
  }
}

template <> struct BigBfly<6> { static B2C_MATH void run(cplx* v) { bfly_pfa<2, 3>(v); } };
template <> struct BigBfly<10> { static B2C_MATH void run(cplx* v) { bfly_pfa<2, 5>(v); } };
template <> struct BigBfly<12> { static B2C_MATH void run(cplx* v) { bfly_pfa<3, 4>(v); } };
template <> struct BigBfly<14> { static B2C_MATH void run(cplx* v) { bfly_pfa<2, 7>(v); } };
template <> struct BigBfly<9> {
  static B2C_MATH void run(cplx* v) {
    const cplx w1 = make_double2(0.76604444311897803520, -0.64278760968653932632);
    const cplx w2 = make_double2(0.17364817766693034885, -0.98480775301220805937);
    const cplx w4 = make_double2(-0.93969262078590838405, -0.34202014332566873304);
    cplx a0[3] = {v[0], v[3], v[6]};
    cplx a1[3] = {v[1], v[4], v[7]};
    cplx a2[3] = {v[2], v[5], v[8]};
    Bfly<3>::run(a0); Bfly<3>::run(a1); Bfly<3>::run(a2);
    a1[1] = cmul(a1[1], w1); a1[2] = cmul(a1[2], w2);
    a2[1] = cmul(a2[1], w2); a2[2] = cmul(a2[2], w4);
#pragma unroll
    for (int k1 = 0; k1 < 3; ++k1) {
      cplx s[3] = {a0[k1], a1[k1], a2[k1]};
      Bfly<3>::run(s);
      v[k1] = s[0]; v[k1 + 3] = s[1]; v[k1 + 6] = s[2];
    }
  }
};
template <> struct BigBfly<16> {   // 4 x 4 Cooley-Tukey, twiddles w16^(n1*k2)
  static B2C_MATH void run(cplx* v) {
    const double c1 = 0.92387953251128675613, s1 = 0.38268343236508977173, h = 0.70710678118654752440;
    cplx a[4][4];   // a[n1][k2]: DFT4 over n2 of v[n1 + 4*n2]
#pragma unroll
    for (int n1 = 0; n1 < 4; ++n1) {
      cplx t[4] = {v[n1], v[n1 + 4], v[n1 + 8], v[n1 + 12]};
      Bfly<4>::run(t);
#pragma unroll
      for (int k2 = 0; k2 < 4; ++k2) a[n1][k2] = t[k2];
    }
    // twiddle a[n1][k2] *= w16^(n1*k2)
    const cplx w1 = make_double2(c1, -s1), w2 = make_double2(h, -h), w3 = make_double2(s1, -c1);
    const cplx w6 = make_double2(-h, -h), w9 = make_double2(-c1, s1);
    a[1][1] = cmul(a[1][1], w1); a[1][2] = cmul(a[1][2], w2); a[1][3] = cmul(a[1][3], w3);
    a[2][1] = cmul(a[2][1], w2); a[2][2] = mul_mi(a[2][2]);   a[2][3] = cmul(a[2][3], w6);
    a[3][1] = cmul(a[3][1], w3); a[3][2] = cmul(a[3][2], w6); a[3][3] = cmul(a[3][3], w9);
#pragma unroll
    for (int k2 = 0; k2 < 4; ++k2) {
      cplx s[4] = {a[0][k2], a[1][k2], a[2][k2], a[3][k2]};
      Bfly<4>::run(s);
#pragma unroll
      for (int k1 = 0; k1 < 4; ++k1) v[k2 + 4 * k1] = s[k1];
    }
  }
};

template <> struct BigBfly<18> { static B2C_MATH void run(cplx* v) { bfly_pfa<2, 9>(v); } };
template <> struct BigBfly<20> { static B2C_MATH void run(cplx* v) { bfly_pfa<4, 5>(v); } };

// One radix-R stage over `nlines` independent lines living in one work buffer.
//   element k of line l is at  W[l*ls + k*es]
//   P    : transform length;  n : current block length (n % R == 0);  m = n / R
//   tw   : tw[k] = exp(-2*pi*i*k/P), k in [0,P)
// Forward (INV=false): DIF -- butterfly, then multiply output q by w_n^(j*q).
// Backward (INV=true) : DIT mirror -- multiply input q by conj(w_n^(j*q)), then inverse butterfly.
// The inverse butterfly is the forward one with re/im swapped on the way in and out.
// `line_fastest` chooses which task index varies fastest across threads so that the unit-stride
// dimension is the one adjacent threads walk (rows: the element index; columns: the line index).
// n / d for 0 <= n < 2^22, 1 <= d: float estimate + one correction step (exact), ~6 instructions
// instead of the ~25 of an integer division.
__device__ __forceinline__ int fast_div(int n, int d, float inv_d) {
  int q = __float2int_rz(__int2float_rz(n) * inv_d);
  const int r = n - q * d;
  if (r < 0) --q;
  else if (r >= d) ++q;
  return q;
}

template <int R, bool INV>
__device__ __forceinline__ void fft_stage(cplx* W, int es, int ls, int nlines, int P, int n,
                                          const cplx* __restrict__ tw, bool line_fastest) {
  const int m = n / R;
  const int per_line = P / R;
  const int total = nlines * per_line;
  const int tstep = P / n;
  const float inv_lines = 1.0f / (float)nlines, inv_per = 1.0f / (float)per_line, inv_m = 1.0f / (float)m;
  for (int task = threadIdx.x; task < total; task += blockDim.x) {
    int line, t;
    if (line_fastest) { t = fast_div(task, nlines, inv_lines); line = task - t * nlines; }
    else { line = fast_div(task, per_line, inv_per); t = task - line * per_line; }
    const int b = fast_div(t, m, inv_m);
    const int j = t - b * m;
    cplx* p = W + (line * ls + (b * n + j) * es);   // 32-bit offsets: a work buffer has < 2^24 elements
    const int stride = m * es;
    cplx v[R];
#pragma unroll
    for (int q = 0; q < R; ++q) v[q] = p[q * stride];
    if (INV) {
      if (j != 0) {
#pragma unroll
        for (int q = 1; q < R; ++q) v[q] = cmulc(v[q], __ldg(&tw[j * q * tstep]));
      }
#pragma unroll
      for (int q = 0; q < R; ++q) v[q] = cswap(v[q]);
      BigBfly<R>::run(v);
#pragma unroll
      for (int q = 0; q < R; ++q) v[q] = cswap(v[q]);
    } else {
      BigBfly<R>::run(v);
      if (j != 0) {
#pragma unroll
        for (int q = 1; q < R; ++q) v[q] = cmul(v[q], __ldg(&tw[j * q * tstep]));
      }
    }
#pragma unroll
    for (int q = 0; q < R; ++q) p[q * stride] = v[q];
  }
}

template <bool INV>
__device__ __forceinline__ void fft_stage_dispatch(int R, cplx* W, int es, int ls, int nlines, int P,
                                                   int n, const cplx* __restrict__ tw, bool lf) {
  switch (R) {
    case 2: fft_stage<2, INV>(W, es, ls, nlines, P, n, tw, lf); break;
    case 3: fft_stage<3, INV>(W, es, ls, nlines, P, n, tw, lf); break;
    case 4: fft_stage<4, INV>(W, es, ls, nlines, P, n, tw, lf); break;
    case 5: fft_stage<5, INV>(W, es, ls, nlines, P, n, tw, lf); break;
    case 6: fft_stage<6, INV>(W, es, ls, nlines, P, n, tw, lf); break;
    case 7: fft_stage<7, INV>(W, es, ls, nlines, P, n, tw, lf); break;
    case 8: fft_stage<8, INV>(W, es, ls, nlines, P, n, tw, lf); break;
    case 9: fft_stage<9, INV>(W, es, ls, nlines, P, n, tw, lf); break;
    case 10: fft_stage<10, INV>(W, es, ls, nlines, P, n, tw, lf); break;
    case 12: fft_stage<12, INV>(W, es, ls, nlines, P, n, tw, lf); break;
  }
}

struct AxisDev {
  int P;
  int nstages;
  int radix[16];
  const cplx* tw;   // exp(-2*pi*i*k/P)
  const int* neg;   // neg[pos] = position holding frequency (-freq(pos)) mod P   (scrambled order)
  const int* posf;  // posf[k]  = position holding frequency k
};

// Forward transform of `nlines` lines (starting at W, i.e. the caller offsets W to the first line).
// Ends with a __syncthreads().
__device__ __forceinline__ void fft_forward(const AxisDev& ax, cplx* W, int es, int ls, int nlines, bool lf) {
  int n = ax.P;
  for (int s = 0; s < ax.nstages; ++s) {
    const int R = ax.radix[s];
    // row-like axis (es == 1): with enough lines, adjacent lanes walk adjacent LINES (stride ls, odd
    // by construction -> no bank conflicts) so that a warp shares one butterfly index and its
    // twiddles are a single broadcast read; with few lines (1-D signals) lanes walk the butterfly index
    const bool stage_lf = lf || nlines >= 16 || (n / R < 8 && nlines >= 8);
    fft_stage_dispatch<false>(R, W, es, ls, nlines, ax.P, n, ax.tw, stage_lf);
    n /= R;
    __syncthreads();
  }
}

// Backward (unnormalised) transform: the forward stages undone in reverse order.
__device__ __forceinline__ void fft_backward(const AxisDev& ax, cplx* W, int es, int ls, int nlines, bool lf) {
  int n = 1;
  for (int s = ax.nstages - 1; s >= 0; --s) {
    const int R = ax.radix[s];
    n *= R;
    const bool stage_lf = lf || nlines >= 16 || (n / R < 8 && nlines >= 8);
    fft_stage_dispatch<true>(R, W, es, ls, nlines, ax.P, n, ax.tw, stage_lf);
    __syncthreads();
  }
}


}  // namespace b2c
