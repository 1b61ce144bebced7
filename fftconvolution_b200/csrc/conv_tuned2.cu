// Family of tuned fused convolution kernels for square periods P x P with P = R1 * R2 (R1 even).
// Period 108 itself (the headline shape) runs the five-pass prime-factor kernel of conv_pfa108.cu; the
// (12, 9) member here is kept for A/B measurements (B2C_P108=t2).  Seven in-place passes over one shared
// buffer Z[P/2][LD]:
//   1 vertical fwd radix-R1 (global -> Z; image columns c and c+P/2 ride one complex transform)
//   2 vertical fwd radix-R2
//   3 horizontal fwd radix-R1 with Hermitian unpack (P/2+1 vertical frequencies)
//   4 horizontal radix-R2 fwd * kernel spectrum * radix-R2 bwd
//   5 horizontal bwd radix-R1 with repack
//   6 vertical bwd radix-R2
//   7 vertical bwd radix-R1, crop, Z -> global
// Replaces fftw_circular_convolution + crop (reference convolution_fftw.h:169-284) for shared-kernel
// plans whose two periods are equal and in the menu at the bottom of this file.
//
// Scrambled order after the (R1, R2) decimation in frequency: position pos = R2*p + s holds
// frequency p + R1*s.  The Hermitian partner of pos is P + R2 - 1 - pos for p >= 1 and R2 - s for
// p == 0; frequencies 0 and P/2 are their own partners and park their second half in two spare
// columns (P, P+1).
#include "conv_params.h"
#include "launch_cache.h"

namespace b2c {

extern __shared__ __align__(16) unsigned char b2c_smem_t2[];

namespace t2 {

struct Params {
  int hs, ws, hd, wd, off_h, off_w;
  // periodic-extension form of CIRCULAR_SAME (planner.h): the transform input has hs_in x ws_in positions,
  // position (r, c) holds source sample ((r - sh_h) mod hs, (c - sh_w) mod ws); hs, ws, 0, 0 otherwise
  int hs_in, ws_in, sh_h, sh_w;
  const double* src;
  double* dst;
  long n_images;
  const cplx* kspec;  // [P][KLD]: (scrambled horizontal position, half-spectrum vertical frequency kk)
  const cplx* tw;     // exp(-2 pi i k / P)
  // per-image kernels (PERIMG instantiation): kernel taps [n][hk][wk] with hk, wk <= P, and one spectrum slot of
  // P * KLD values per CTA in `kscratch` (L2-resident), rewritten for every image
  const double* ker;
  int hk, wk;
  cplx* kscratch;
};

template <int R> using Big = BigBfly<R>;   // forward DFT of R points in registers (fft_device.cuh)

// Predicated global accesses of passes 1 and 7.  B2C_T2_ASM_PRED=1: one predicated instruction each (inline PTX)
// instead of what the compiler makes of a conditional expression (select / branch per element).
#ifndef B2C_T2_ASM_PRED
#define B2C_T2_ASM_PRED 0
#endif
__device__ __forceinline__ double ldg_if(const double* ptr, bool on) {
  double v;
  asm volatile("{\n .reg .pred p;\n setp.ne.u32 p, %2, 0;\n mov.f64 %0, 0d0000000000000000;\n @p ld.global.nc.f64 %0, [%1];\n}"
               : "=d"(v) : "l"(ptr), "r"((unsigned)on));
  return v;
}
__device__ __forceinline__ void stg_if(double* ptr, double v, bool on) {
  asm volatile("{\n .reg .pred p;\n setp.ne.u32 p, %2, 0;\n @p st.global.f64 [%0], %1;\n}" ::"l"(ptr), "d"(v), "r"((unsigned)on) : "memory");
}
#if B2C_T2_ASM_PRED
#define B2C_T2_LDG_IF(ptr, on) ldg_if(ptr, on)
#define B2C_T2_STG_IF(ptr, val, on) stg_if(ptr, val, on)
#else
#define B2C_T2_LDG_IF(ptr, on) ((on) ? __ldg(ptr) : 0.0)
#define B2C_T2_STG_IF(ptr, val, on) do { if (on) *(ptr) = (val); } while (0)
#endif

#ifndef B2C_T2_TP
#define B2C_T2_TP 0    // A/B builds: period whose CTA size is overridden ...
#define B2C_T2_TT 0    // ... and the size
#endif
#ifndef B2C_T2_TP_MINB
#define B2C_T2_TP_MINB 0   // A/B builds: resident CTAs the registers of that period's kernel are sized for (0: by formula)
#endif
#ifndef B2C_T2_TP_TWMEM
#define B2C_T2_TP_TWMEM 0  // A/B builds: that period reads its twiddles from memory instead of keeping them in registers
#endif
// CTA size by period: 64 threads up to 48, 128 up to 80, 256 above (measured: P = 40: 74 -> 113 M conv/s, P = 80: 23.3 -> 28.5
// against 256 threads), except where that size leaves a pass with a nearly empty last round of tasks (a pass hands
// out lines x radix tasks in rounds of THREADS): measured with tools/gpu_periods.sh, ns per convolution
//   P = 42: 11.5 -> 10.8 (160)   50: 16.7 -> 16.2 (160)   60: 22.3 -> 19.9 (96)   84: 52.3 -> 44.5 (160)
//   100: 72.6 -> 69.1 (192)   126: 142 -> 134 (224);   no gain at 54, 64, 72, 80, 90, 96, 112, 120.
constexpr int threads_of(int P) {
  return P == 42 ? 160 : P == 50 ? 160 : P == 60 ? 96 : P == 84 ? 160 : P == 100 ? 192 : P == 126 ? 224
       : P <= 48 ? 64 : P <= 80 ? 128 : 256;
}

template <int R1, int R2>
struct Geo {
  static constexpr int P = R1 * R2;
  static constexpr int HALF = P / 2;
  static constexpr int H1 = R1 / 2;
  static constexpr int NK = P / 2 + 1;
  static constexpr int LD = ((P + 2) | 1) + ((((P + 2) | 1) % 8 == 1) ? 2 : 0);   // odd, not == 1 (mod 8)
  static constexpr int KLD = (NK + 7) / 8 * 8;
  static constexpr int BPAD = KLD;
  // CTA size follows the work per pass (threads_of above)
  static constexpr int THREADS = P == B2C_T2_TP ? B2C_T2_TT : threads_of(P);
  static constexpr int A_LINES = THREADS / R2;
  static constexpr int A_ROUNDS = (NK + A_LINES - 1) / A_LINES;
  static constexpr int N0 = R2 / 2 + 1;          // half-spectrum columns in the p == 0 family
  static constexpr bool TW_IN_REGS = R1 <= 12 && !(B2C_T2_TP_TWMEM && P == B2C_T2_TP);   // R1-1 complex twiddles per thread
  static_assert(R1 % 2 == 0, "the column packing pairs c with c + P/2 = c + R2 * R1/2");
  static __host__ __device__ constexpr int freq_of_pos(int pos) { return pos / R2 + R1 * (pos % R2); }
  static __host__ __device__ constexpr int rep_of(int kk) { return kk < N0 ? kk : R2 + (kk - N0); }
  static __host__ __device__ constexpr bool self_of(int kk) {
    return kk == 0 || (kk < N0 ? (R2 % 2 == 0 && kk == R2 / 2) : (P + R2 - 1 - rep_of(kk) == rep_of(kk)));
  }
  static __host__ __device__ constexpr int par_of(int kk) {
    return kk == 0 ? P : (self_of(kk) ? P + 1 : (kk < N0 ? R2 - kk : P + R2 - 1 - rep_of(kk)));
  }
  static constexpr size_t smem_bytes = sizeof(cplx) * (size_t)HALF * LD;
  // resident CTAs the register allocation is sized for: 128 registers per thread, or what shared memory allows
  static constexpr int BY_SMEM = (int)((227 * 1024) / (smem_bytes + 1024)), BY_REGS = 512 / THREADS;
  static constexpr int MIN_CTAS = (B2C_T2_TP_MINB && P == B2C_T2_TP) ? B2C_T2_TP_MINB : BY_SMEM < BY_REGS ? (BY_SMEM < 1 ? 1 : BY_SMEM) : BY_REGS;
};

#define B2C_T2_PERIMG 0
#define B2C_T2_KERNEL conv_kernel
#include "conv_tuned2_kernel.inc"
#undef B2C_T2_PERIMG
#undef B2C_T2_KERNEL
#define B2C_T2_PERIMG 1
#define B2C_T2_KERNEL conv_kernel_perimage
#include "conv_tuned2_kernel.inc"
#undef B2C_T2_PERIMG
#undef B2C_T2_KERNEL

// Kernel spectrum K[pos][kk] = 0.5/(P*P) * DFT2(kernel)[u(kk)][v(pos)], straight from the taps, as a separable sum:
// T1[b] = sum_a tap[a][b] w^(u a), then K[pos][kk] = sum_b T1[b] w^(v b).  One CTA per spectrum column kk: it builds
// the T1 row of its own vertical frequency u(kk) (threads over the kernel columns b, taps read coalesced) and then all
// P positions of that column (threads over pos).  No work is repeated between CTAs, so the cost is hke*wke + P*wke
// complex FMAs per CTA whatever the kernel size (round 2 history: one 512-thread CTA for everything took 34 us at
// 15 x 15 taps and 450 us at 100 x 100; sums run in the same order as there, so the spectra are bit-identical).
// (mh, mw): the moduli the taps fold by (convolution_fftw.h:185-187: taps congruent modulo the period add up) -- the
// period, or the source extents for a CIRCULAR_SAME kernel larger than its source computed by periodic extension
// (planner.cpp: make_plan_circular_ext).
constexpr int kKspecThreads = 256;
template <int R1, int R2>
__global__ void __launch_bounds__(kKspecThreads) kspec_kernel(const double* __restrict__ ker, int hk, int wk, int mh, int mw,
                                                              const cplx* __restrict__ tw, cplx* __restrict__ out) {
  typedef Geo<R1, R2> G;
  constexpr int P = G::P, KLD = G::KLD;
  __shared__ cplx roots[P];
  __shared__ cplx T1[P];
  const int hke = min(hk, mh), wke = min(wk, mw);
  const int kk = blockIdx.x;
  const int u = G::freq_of_pos(G::rep_of(kk));
  for (int i = threadIdx.x; i < P; i += kKspecThreads) roots[i] = tw[i];
  __syncthreads();
  for (int b = threadIdx.x; b < wke; b += kKspecThreads) {
    cplx acc = make_double2(0.0, 0.0);
    int ua = 0;   // (u * a) % P
    for (int a = 0; a < hke; ++a) {
      double tap = 0.0;
      for (int aa = a; aa < hk; aa += mh)
        for (int bb = b; bb < wk; bb += mw) tap += ker[(size_t)aa * wk + bb];
      const cplx w = roots[ua];
      acc.x = fma(tap, w.x, acc.x);
      acc.y = fma(tap, w.y, acc.y);
      ua += u; if (ua >= P) ua -= P;
    }
    T1[b] = acc;
  }
  __syncthreads();
  const double scale = 0.5 / ((double)P * (double)P);
  for (int pos = threadIdx.x; pos < P; pos += kKspecThreads) {
    const int v = G::freq_of_pos(pos);
    cplx acc = make_double2(0.0, 0.0);
    int vb = 0;   // (v * b) % P
    for (int b = 0; b < wke; ++b) {
      const cplx a = T1[b];
      const cplx w = roots[vb];
      acc.x += a.x * w.x - a.y * w.y;
      acc.y += a.x * w.y + a.y * w.x;
      vb += v; if (vb >= P) vb -= P;
    }
    out[pos * KLD + kk] = make_double2(acc.x * scale, acc.y * scale);
  }
}

template <int R1, int R2>
static cudaError_t launch_kspec(const double* d_ker, int hk, int wk, int mh, int mw, const cplx* tw, cplx* out, cudaStream_t st) {
  typedef Geo<R1, R2> G;
  if (mh < 1 || mh > G::P) mh = G::P;
  if (mw < 1 || mw > G::P) mw = G::P;
  kspec_kernel<R1, R2><<<G::NK, kKspecThreads, 0, st>>>(d_ker, hk, wk, mh, mw, tw, out);
  return cudaGetLastError();
}

template <int R1, int R2>
cudaError_t run(const Params& p, const double* d_ker, int hk, int wk, int mh, int mw, int sms, cudaStream_t st, int* launches) {
  typedef Geo<R1, R2> G;
  if (d_ker) {   // nullptr: the spectrum of the previous call is still valid (single-call kernel cache)
    cudaError_t e = launch_kspec<R1, R2>(d_ker, hk, wk, mh, mw, p.tw, const_cast<cplx*>(p.kspec), st);
    if (e != cudaSuccess) return e;
    ++*launches;
  }
  int occ = 0;
  occ = cached_occupancy(reinterpret_cast<const void*>(conv_kernel<R1, R2>), G::THREADS, G::smem_bytes);
  if (occ < 1) return cudaErrorLaunchOutOfResources;
  long grid = (long)sms * occ;
  if (grid > p.n_images) grid = p.n_images;
  conv_kernel<R1, R2><<<(int)grid, G::THREADS, G::smem_bytes, st>>>(p);
  ++*launches;
  return cudaGetLastError();
}

// per-image kernels: *grid_out (when non-null) only asks for the grid this launch would use, i.e. how many
// spectrum slots `kscratch` needs
template <int R1, int R2>
cudaError_t run_perimage(const Params& p, int sms, cudaStream_t st, int* launches, long* grid_out) {
  typedef Geo<R1, R2> G;
  int occ = 0;
  occ = cached_occupancy(reinterpret_cast<const void*>(conv_kernel_perimage<R1, R2>), G::THREADS, G::smem_bytes);
  if (occ < 1) return cudaErrorLaunchOutOfResources;
  long grid = (long)sms * occ;
  if (grid_out) { *grid_out = grid; return cudaSuccess; }
  if (grid > p.n_images) grid = p.n_images;
  conv_kernel_perimage<R1, R2><<<(int)grid, G::THREADS, G::smem_bytes, st>>>(p);
  ++*launches;
  return cudaGetLastError();
}

}  // namespace t2

// ---- the menu: square periods with a tuned kernel --------------------------------------------------
#define B2C_T2_MENU(X)                                                                                         \
  X(8, 2) X(10, 2) X(8, 3) X(4, 7) X(8, 4) X(6, 6) X(8, 5) X(6, 7) X(8, 6) X(10, 5) X(6, 9) X(8, 7) X(10, 6) X(8, 8) \
  X(10, 7) X(8, 9) X(10, 8) X(12, 7) X(10, 9) X(12, 8) X(10, 10) X(12, 9) X(14, 8) X(12, 10) X(14, 9) X(16, 8) \
  X(14, 10) X(16, 9) X(16, 10)

bool tuned2_has(int P) {
  switch (P) {
#define X(a, b) case (a) * (b):
    B2C_T2_MENU(X)
#undef X
    return true;
  }
  return false;
}

size_t tuned2_kspec_bytes(int P) { return sizeof(cplx) * (size_t)P * (size_t)((P / 2 + 1 + 7) / 8 * 8); }

// kernel spectrum + convolution for a square period P from the menu; *launches += 2
cudaError_t launch_tuned2(int P, int hs, int ws, int hk, int wk, int hd, int wd, int off_h, int off_w,
                          int hs_in, int ws_in, int sh_h, int sh_w, const double* d_src, const double* d_ker, double* d_dst, long n, cplx* kspec,
                          const cplx* tw, int sms, cudaStream_t st, int* launches, int kmod_h, int kmod_w) {
  t2::Params p;
  p.hs = hs; p.ws = ws; p.hd = hd; p.wd = wd; p.off_h = off_h; p.off_w = off_w;
  p.hs_in = hs_in; p.ws_in = ws_in; p.sh_h = sh_h; p.sh_w = sh_w;
  p.src = d_src; p.dst = d_dst; p.n_images = n; p.kspec = kspec; p.tw = tw;
  p.ker = nullptr; p.hk = hk; p.wk = wk; p.kscratch = nullptr;
  switch (P) {
#define X(a, b) case (a) * (b): return t2::run<a, b>(p, d_ker, hk, wk, kmod_h, kmod_w, sms, st, launches);
    B2C_T2_MENU(X)
#undef X
  }
  return cudaErrorInvalidValue;
}

// Per-image kernels on a square menu period (kernel extents <= P): `kscratch` holds one spectrum slot of
// tuned2_kspec_bytes(P) per CTA of the launch; with d_src == nullptr only *slots is filled in (nothing runs).
cudaError_t launch_tuned2_perimage(int P, int hs, int ws, int hk, int wk, int hd, int wd, int off_h, int off_w,
                                   const double* d_src, const double* d_ker, double* d_dst, long n, cplx* kscratch,
                                   const cplx* tw, int sms, cudaStream_t st, int* launches, long* slots) {
  t2::Params p;
  p.hs = hs; p.ws = ws; p.hd = hd; p.wd = wd; p.off_h = off_h; p.off_w = off_w;
  p.hs_in = hs; p.ws_in = ws; p.sh_h = 0; p.sh_w = 0;
  p.src = d_src; p.dst = d_dst; p.n_images = n; p.kspec = nullptr; p.tw = tw;
  p.ker = d_ker; p.hk = hk; p.wk = wk; p.kscratch = kscratch;
  switch (P) {
#define X(a, b) case (a) * (b): return t2::run_perimage<a, b>(p, sms, st, launches, d_src ? nullptr : slots);
    B2C_T2_MENU(X)
#undef X
  }
  return cudaErrorInvalidValue;
}

}  // namespace b2c
