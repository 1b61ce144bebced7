// Tuned fused FFT convolution for 1-D signals (1 x N rows; N x 1 columns are planned as rows) with a
// shared kernel -- the long-kernel half of BASELINE config 4, where the direct kernel loses.
// Replaces fftw_circular_convolution + crop (reference convolution_fftw.h:169-284) for plans whose
// horizontal period is on the menu below and whose vertical period is 1.
//
// One CTA transforms TWO signals at a time as one complex sequence z = a + i*b (the convolution is
// linear, so Re/Im of the result are the two convolutions).  The period P = R0*R1*R2(*R3) is a
// compile-time radix schedule of big butterflies (8..20 points in registers), one butterfly per
// thread and pass, in place in one shared buffer:
//   pass 0      : global -> registers (zero padding in registers, coalesced) -> butterfly -> shared
//   passes 1..  : shared -> butterfly -> shared
//   last pass   : forward butterfly, * cached kernel spectrum (1/P folded in), backward butterfly
//   ... mirrored backward passes; the last one crops and writes both signals to global.
// 2*S - 1 trips through shared memory for S passes (the any-size kernel needs 2*S' + 1 with S' = 5
// radix-3/8 passes at P = 4608, most of its instructions index arithmetic).  Forward passes are
// decimation in frequency, backward passes the mirrored decimation in time: the spectrum stays in
// scrambled order and the kernel spectrum is produced in the same order by the same forward code.
// Positions are padded by one slot per LAST-radix block so that the last pass (a thread owns a
// contiguous block) is free of bank conflicts.
#include "conv_params.h"
#include "launch_cache.h"

namespace b2c {

extern __shared__ __align__(16) unsigned char b2c_smem_t1[];

namespace t1 {

#ifndef B2C_T1_POW
#define B2C_T1_POW 1
#endif
#ifndef B2C_T1_PREFETCH
#define B2C_T1_PREFETCH 1
#endif

struct Params {
  int ns, nd, off;          // signal length, result length, crop offset
  int n_in, shift;          // input positions and periodic-extension shift: position q holds sample (q - shift) mod ns
  const double* src;        // [n][ns]
  double* dst;              // [n][nd]
  long n_signals;
  const cplx* kspec;        // [LAST][P / LAST]: spectrum at scrambled position b * LAST + q in entry [q][b], scaled by 1/P
  const cplx* tw;           // exp(-2 pi i k / P)
  const cplx* twtab;        // per-pass twiddle tables [q-1][j] (lanes read consecutive entries), see Geo::OFF*
};

template <int P, int R0, int R1, int R2, int R3>
struct Geo {
  static_assert(R0 * R1 * R2 * R3 == P, "radix schedule");
  static constexpr int NS = R3 > 1 ? 4 : 3;
  static constexpr int LAST = R3 > 1 ? R3 : R2;
  static constexpr int cmax(int a, int b) { return a > b ? a : b; }
  static constexpr int cmin(int a, int b) { return a < b ? a : b; }
  static constexpr int RMIN = cmin(cmin(R0, R1), cmin(R2, R3 > 1 ? R3 : R2));
  static constexpr int TASKS_MAX = P / RMIN;
  // one task per thread in the widest pass when that fits in 512 threads, else two rounds
  static constexpr int THREADS = TASKS_MAX <= 512 ? (TASKS_MAX + 31) / 32 * 32 : (TASKS_MAX / 2 + 31) / 32 * 32;
  static constexpr int SLOTS = P + P / LAST;
  static constexpr size_t smem_bytes = sizeof(cplx) * (size_t)SLOTS;
  static __host__ __device__ constexpr int pad(int p) { return p + p / LAST; }
  static constexpr bool ONE_ROUND = P / R0 <= THREADS;   // first pass: one butterfly per thread
  static constexpr int MIN_CTAS = 2 * smem_bytes <= 225 * 1024 ? 2 : 1;   // register budget follows the CTAs that fit
  // Twiddle tables, one per pass that has twiddles: entry (q-1) * M + j = w_N^(j*q).  Reading tw[j*q*TSTEP]
  // instead costs one 32-byte sector per lane (ncu: L1TEX 83 % busy, long-scoreboard the top stall).
  static constexpr int OFF0 = 0;
  static constexpr int OFF1 = OFF0 + (R0 - 1) * (P / R0);
  static constexpr int OFF2 = OFF1 + (R1 - 1) * (P / R0 / R1);
  static constexpr int TABN = OFF2 + (NS == 4 ? (R2 - 1) * (P / R0 / R1 / R2) : 0);
};

// One pass.  N: block length of this pass, R: radix, M = N / R.  Element q of butterfly (b, j) sits
// at position b*N + j + M*q; twiddle of output q (forward) / input q (backward) is w_N^(j*q).
//   SRC_GLOBAL : inputs come from the two signals in global memory (first forward pass)
//   DST_GLOBAL : outputs go to the two results in global memory, cropped (last backward pass)
//   MULK       : M == 1: forward butterfly, multiply by the kernel spectrum, backward butterfly
template <class G, int P, int N, int R, bool INV, bool SRC_GLOBAL, bool DST_GLOBAL, bool MULK, int TOFF>
__device__ __forceinline__ void pass(cplx* W, const Params& p, const double* a, const double* b2, double* da, double* db,
                                     cplx wj = make_double2(1.0, 0.0), double bs = 1.0) {   // bs: factor on signal b (see first_pass_pair)
  constexpr int M = N / R;
  // first / last pass with one butterfly per thread: the thread's j never changes, its root w_P^j lives in
  // registers and the R-1 twiddles are its powers (17 dependent multiplications cost less than 17 table
  // reads that together move as many bytes through L2 as the signals themselves)
  constexpr bool POW = B2C_T1_POW && N == P && (P / R <= G::THREADS) && R <= 16;   // radix 18 and 20 spill with it
  const cplx* tab = p.twtab + (TOFF >= 0 ? TOFF : 0);
  for (int task = threadIdx.x; task < P / R; task += G::THREADS) {
    const int blk = task / M, j = task - blk * M;
    const int base = blk * N + j;
    cplx v[R];
    if (SRC_GLOBAL) {
#pragma unroll
      for (int q = 0; q < R; ++q) {
        const int pos = base + M * q;
        v[q] = make_double2(0.0, 0.0);
        if (pos < p.n_in) {
          int si = pos - p.shift;
          if (si < 0) si += p.ns;
          v[q].x = __ldg(a + si);
          if (b2) v[q].y = bs * __ldg(b2 + si);
        }
      }
      if (p.n_in > P) {   // samples congruent modulo the period add up (a kernel longer than the period,
                          // convolution_fftw.h:185-187); kept out of the loads above so that they stay independent
#pragma unroll
        for (int q = 0; q < R; ++q)
          for (int i = base + M * q + P; i < p.n_in; i += P) {
            v[q].x += __ldg(a + i);
            if (b2) v[q].y += __ldg(b2 + i);
          }
      }
    } else {
#pragma unroll
      for (int q = 0; q < R; ++q) v[q] = W[G::pad(base + M * q)];
    }
    if (MULK) {
      BigBfly<R>::run(v);
#pragma unroll
      for (int q = 0; q < R; ++q) {
        const cplx y = cmul(v[q], __ldg(p.kspec + q * (P / R) + blk));   // [q][block]: lanes read consecutive entries
        v[q] = make_double2(y.y, y.x);          // swapped: backward butterfly = forward on (im, re)
      }
      BigBfly<R>::run(v);
#pragma unroll
      for (int q = 0; q < R; ++q) v[q] = cswap(v[q]);
    } else if (INV) {
      if (M > 1) {
        if (POW) {   // w^(j*q) by repeated multiplication from the thread's own w^j (see conv_kernel)
          cplx cur = wj;
#pragma unroll
          for (int q = 1; q < R; ++q) { v[q] = cmulc(v[q], cur); cur = cmul(cur, wj); }
        } else {
#pragma unroll
          for (int q = 1; q < R; ++q) v[q] = cmulc(v[q], __ldg(tab + (q - 1) * M + j));
        }
      }
#pragma unroll
      for (int q = 0; q < R; ++q) v[q] = cswap(v[q]);
      BigBfly<R>::run(v);
#pragma unroll
      for (int q = 0; q < R; ++q) v[q] = cswap(v[q]);
    } else {
      BigBfly<R>::run(v);
      if (M > 1) {
        if (POW) {
          cplx cur = wj;
#pragma unroll
          for (int q = 1; q < R; ++q) { v[q] = cmul(v[q], cur); cur = cmul(cur, wj); }
        } else {
#pragma unroll
          for (int q = 1; q < R; ++q) v[q] = cmul(v[q], __ldg(tab + (q - 1) * M + j));
        }
      }
    }
    if (DST_GLOBAL) {
#pragma unroll
      for (int q = 0; q < R; ++q) {
        const int d = base + M * q - p.off;
        if (d >= 0 && d < p.nd) {
          da[d] = v[q].x;
          if (db) db[d] = v[q].y * bs;
        }
      }
    } else {
#pragma unroll
      for (int q = 0; q < R; ++q) W[G::pad(base + M * q)] = v[q];
    }
  }
}

// First forward pass of a PAIR of signals with one butterfly per thread.  Two signals ride one complex
// transform; the rounding noise of the larger one would land on the smaller one, so signal b is scaled by the
// power of two `beta` that equalises the two maxima (exact; undone when the result is stored) -- the reference
// transforms every signal on its own (convolution_fftw.h:190).  All threads of the CTA must call this
// (one __syncthreads inside); returns beta.
template <class G, int P, int R>
__device__ __forceinline__ double first_pass_pair(cplx* W, const Params& p, const double* a, const double* b2, cplx wj,
                                                  unsigned* red) {
  constexpr int M = P / R;
  static_assert(M <= G::THREADS, "one butterfly per thread");
  const bool active = threadIdx.x < M;
  const int j = active ? threadIdx.x : 0;
  cplx v[R];
#pragma unroll
  for (int q = 0; q < R; ++q) {
    const int pos = j + M * q;
    v[q] = make_double2(0.0, 0.0);
    if (active && pos < p.n_in) {
      int si = pos - p.shift;
      if (si < 0) si += p.ns;
      v[q].x = __ldg(a + si);
      v[q].y = __ldg(b2 + si);
    }
  }
  double ma = 0.0, mb = 0.0;
#pragma unroll
  for (int q = 0; q < R; ++q) { ma = fmax(ma, fabs(v[q].x)); mb = fmax(mb, fabs(v[q].y)); }
  const unsigned wa = __reduce_max_sync(0xffffffffu, (unsigned)__double2hiint(ma)) >> 20;   // exponent words
  const unsigned wb = __reduce_max_sync(0xffffffffu, (unsigned)__double2hiint(mb)) >> 20;
  const int warp = threadIdx.x >> 5;
  constexpr int NW = G::THREADS / 32;
  if ((threadIdx.x & 31) == 0) { red[warp] = wa; red[NW + warp] = wb; }
  __syncthreads();
  unsigned ea = 0, eb = 0;
#pragma unroll
  for (int w = 0; w < NW; ++w) { ea = max(ea, red[w]); eb = max(eb, red[NW + w]); }
  const bool usable = ea > 0 && eb > 0 && ea < 2047 && eb < 2047;   // zero / subnormal / inf / nan: no scaling
  const double beta = usable ? __hiloint2double((int)((1023u + ea - eb) << 20), 0) : 1.0;
  if (active) {
#pragma unroll
    for (int q = 0; q < R; ++q) v[q].y *= beta;
    BigBfly<R>::run(v);
    constexpr bool POW = B2C_T1_POW && R <= 16;
    if (POW) {
      cplx cur = wj;
#pragma unroll
      for (int q = 1; q < R; ++q) { v[q] = cmul(v[q], cur); cur = cmul(cur, wj); }
    } else {
#pragma unroll
      for (int q = 1; q < R; ++q) v[q] = cmul(v[q], __ldg(p.twtab + G::OFF0 + (q - 1) * M + j));
    }
#pragma unroll
    for (int q = 0; q < R; ++q) W[G::pad(j + M * q)] = v[q];
  }
  return beta;
}

template <int P, int R0, int R1, int R2, int R3>
__device__ __forceinline__ void forward_all_but_last(cplx* W, const Params& p, const double* a, const double* b2, cplx wj,
                                                     bool first_done = false, double bs = 1.0) {
  typedef Geo<P, R0, R1, R2, R3> G;
  if (!first_done) pass<G, P, P, R0, false, true, false, false, G::OFF0>(W, p, a, b2, nullptr, nullptr, wj, bs);
  __syncthreads();
  if (G::NS == 4) {
    pass<G, P, P / R0, R1, false, false, false, false, G::OFF1>(W, p, nullptr, nullptr, nullptr, nullptr);
    __syncthreads();
    pass<G, P, P / R0 / R1, R2, false, false, false, false, G::OFF2>(W, p, nullptr, nullptr, nullptr, nullptr);
    __syncthreads();
  } else {
    pass<G, P, P / R0, R1, false, false, false, false, G::OFF1>(W, p, nullptr, nullptr, nullptr, nullptr);
    __syncthreads();
  }
}

template <int P, int R0, int R1, int R2, int R3>
__global__ void __launch_bounds__((Geo<P, R0, R1, R2, R3>::THREADS), (Geo<P, R0, R1, R2, R3>::MIN_CTAS)) conv_kernel(const __grid_constant__ Params p) {
  typedef Geo<P, R0, R1, R2, R3> G;
  cplx* W = reinterpret_cast<cplx*>(b2c_smem_t1);
  __shared__ unsigned red[2 * (G::THREADS / 32)];
  const long pairs = (p.n_signals + 1) / 2;
  const cplx wj = __ldg(p.tw + (threadIdx.x < P / R0 ? threadIdx.x : 0));   // w_P^j of this thread's first-pass butterfly
  for (long pr = blockIdx.x; pr < pairs; pr += gridDim.x) {
    const double* a = p.src + (size_t)(2 * pr) * p.ns;
    const bool two = 2 * pr + 1 < p.n_signals;
#if B2C_T1_PREFETCH
    if (threadIdx.x == 0 && pr + gridDim.x < pairs) {   // next pair -> L2: its first pass meets L2 latency, not HBM latency
      const long nxt = 2 * (pr + gridDim.x);
      l2_prefetch(p.src + (size_t)nxt * p.ns, (size_t)(nxt + 1 < p.n_signals ? 2 : 1) * p.ns * sizeof(double));
    }
#endif
    const double* b2 = two ? a + p.ns : nullptr;
    double* da = p.dst + (size_t)(2 * pr) * p.nd;
    double* db = two ? da + p.nd : nullptr;
    double beta = 1.0;   // power-of-two factor on signal b that equalises the pair's magnitudes
    if (two) {
      if constexpr (G::ONE_ROUND) {
        beta = first_pass_pair<G, P, R0>(W, p, a, b2, wj, red);
        forward_all_but_last<P, R0, R1, R2, R3>(W, p, a, b2, wj, true);
      } else {
        beta = cta_balance(a, p.ns, b2, p.ns, reinterpret_cast<double*>(W));
        forward_all_but_last<P, R0, R1, R2, R3>(W, p, a, b2, wj, false, beta);
      }
    } else {
      forward_all_but_last<P, R0, R1, R2, R3>(W, p, a, b2, wj);
    }
    pass<G, P, G::LAST, G::LAST, false, false, false, true, -1>(W, p, nullptr, nullptr, nullptr, nullptr);
    __syncthreads();
    if (G::NS == 4) {
      pass<G, P, P / R0 / R1, R2, true, false, false, false, G::OFF2>(W, p, nullptr, nullptr, nullptr, nullptr);
      __syncthreads();
    }
    pass<G, P, P / R0, R1, true, false, false, false, G::OFF1>(W, p, nullptr, nullptr, nullptr, nullptr);
    __syncthreads();
    pass<G, P, P, R0, true, false, true, false, G::OFF0>(W, p, nullptr, nullptr, da, db, wj, 1.0 / beta);
    __syncthreads();   // the next pair's first pass overwrites W
  }
}

// Kernel spectrum in the scrambled order the product pass reads, scaled by 1/P: the same forward
// passes applied to the zero-padded taps (one CTA, once per batch; the reference recomputes the
// kernel transform for every signal, convolution_fftw.h:191).
template <int P, int R0, int R1, int R2, int R3>
__global__ void __launch_bounds__((Geo<P, R0, R1, R2, R3>::THREADS), 1) kspec_kernel(const __grid_constant__ Params p, cplx* out) {
  typedef Geo<P, R0, R1, R2, R3> G;
  cplx* W = reinterpret_cast<cplx*>(b2c_smem_t1);
  forward_all_but_last<P, R0, R1, R2, R3>(W, p, p.src, nullptr, __ldg(p.tw + (threadIdx.x < P / R0 ? threadIdx.x : 0)));
  pass<G, P, G::LAST, G::LAST, false, false, false, false, -1>(W, p, nullptr, nullptr, nullptr, nullptr);
  __syncthreads();
  const double scale = 1.0 / (double)P;
  for (int i = threadIdx.x; i < P; i += G::THREADS) {
    const cplx v = W[G::pad(i)];
    out[(i % G::LAST) * (P / G::LAST) + i / G::LAST] = make_double2(v.x * scale, v.y * scale);   // position b*LAST + q -> [q][b]
  }
}

// Fill the per-pass twiddle tables from the plan's root table (once per plan).
template <int P, int R0, int R1, int R2, int R3>
__global__ void twtab_kernel(const cplx* __restrict__ tw, cplx* __restrict__ tab) {
  typedef Geo<P, R0, R1, R2, R3> G;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < G::TABN; i += gridDim.x * blockDim.x) {
    int off, m, tstep;   // table of the pass with block length N = P / tstep, M = N / radix
    if (i < G::OFF1) { off = G::OFF0; m = P / R0; tstep = 1; }
    else if (i < G::OFF2) { off = G::OFF1; m = P / R0 / R1; tstep = R0; }
    else { off = G::OFF2; m = P / R0 / R1 / R2; tstep = R0 * R1; }
    const int k = i - off, q = k / m + 1, j = k - (q - 1) * m;
    tab[i] = tw[j * q * tstep];
  }
}

template <int P, int R0, int R1, int R2, int R3>
cudaError_t run(const Params& p, const double* d_ker, int nk, cplx* kspec, int build_table, int sms, cudaStream_t st,
                int* launches) {
  typedef Geo<P, R0, R1, R2, R3> G;
  if (build_table) {
    twtab_kernel<P, R0, R1, R2, R3><<<(G::TABN + 255) / 256, 256, 0, st>>>(p.tw, const_cast<cplx*>(p.twtab));
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    ++*launches;
  }
  auto kk = kspec_kernel<P, R0, R1, R2, R3>;
  auto ck = conv_kernel<P, R0, R1, R2, R3>;
  ensure_dyn_smem(reinterpret_cast<const void*>(kk), G::smem_bytes);
  ensure_dyn_smem(reinterpret_cast<const void*>(ck), G::smem_bytes);
  Params pk = p;
  pk.src = d_ker; pk.ns = nk; pk.n_in = nk; pk.shift = 0; pk.n_signals = 1;
  if (d_ker) {   // nullptr: the spectrum of the previous call is still valid (single-call kernel cache)
    kk<<<1, G::THREADS, G::smem_bytes, st>>>(pk, kspec);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    ++*launches;
  }
  int occ = 0;
  occ = cached_occupancy(reinterpret_cast<const void*>(ck), G::THREADS, G::smem_bytes);
  if (occ < 1) return cudaErrorLaunchOutOfResources;
  const long pairs = (p.n_signals + 1) / 2;
  long grid = (long)sms * occ;
  if (grid > pairs) grid = pairs;
  ck<<<(int)grid, G::THREADS, G::smem_bytes, st>>>(p);
  ++*launches;
  return cudaGetLastError();
}

}  // namespace t1

// ---- the menu: (P, R0, R1, R2, R3); the planner prefers these periods for 1-D signals ------------
#define B2C_T1_MENU(X)                                                                                          \
  X(128, 4, 4, 8, 1) X(192, 4, 6, 8, 1) X(256, 4, 8, 8, 1) X(320, 5, 8, 8, 1) X(384, 6, 8, 8, 1)                  \
  X(512, 8, 8, 8, 1) X(640, 8, 8, 10, 1) X(768, 8, 8, 12, 1) X(1024, 8, 8, 16, 1) X(1280, 8, 10, 16, 1)         \
  X(1536, 8, 12, 16, 1) X(2048, 8, 16, 16, 1) X(2304, 9, 16, 16, 1) X(2560, 10, 16, 16, 1) X(3072, 12, 16, 16, 1) \
  X(4096, 16, 16, 16, 1) X(4608, 18, 16, 16, 1) X(5120, 20, 16, 16, 1) X(6144, 8, 8, 8, 12) X(8192, 8, 8, 8, 16)

bool tuned1d_has(int P) {
  switch (P) {
#define X(p, a, b, c, d) case p:
    B2C_T1_MENU(X)
#undef X
    return true;
  }
  return false;
}

// Scratch of a 1-D plan, in complex elements: the kernel spectrum [P] followed by the twiddle tables (< 3 P).
size_t tuned1d_scratch_elems(int P) { return (size_t)4 * P; }

// (twiddle tables once per plan) + kernel spectrum + convolution for a 1-D period P from the menu
cudaError_t launch_tuned1d(int P, int ns, int nk, int nd, int off, int n_in, int shift, const double* d_src, const double* d_ker,
                           double* d_dst, long n, cplx* scratch, int build_table, const cplx* tw, int sms,
                           cudaStream_t st, int* launches) {
  t1::Params p;
  cplx* kspec = scratch;
  p.ns = ns; p.nd = nd; p.off = off; p.n_in = n_in; p.shift = shift; p.src = d_src; p.dst = d_dst; p.n_signals = n; p.kspec = kspec; p.tw = tw;
  p.twtab = scratch + P;
  switch (P) {
#define X(pp, a, b, c, d) case pp: return t1::run<pp, a, b, c, d>(p, d_ker, nk, kspec, build_table, sms, st, launches);
    B2C_T1_MENU(X)
#undef X
  }
  return cudaErrorInvalidValue;
}

}  // namespace b2c
