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
};

template <int R> using Big = BigBfly<R>;   // forward DFT of R points in registers (fft_device.cuh)

template <int R1, int R2>
struct Geo {
  static constexpr int P = R1 * R2;
  static constexpr int HALF = P / 2;
  static constexpr int H1 = R1 / 2;
  static constexpr int NK = P / 2 + 1;
  static constexpr int LD = ((P + 2) | 1) + ((((P + 2) | 1) % 8 == 1) ? 2 : 0);   // odd, not == 1 (mod 8)
  static constexpr int KLD = (NK + 7) / 8 * 8;
  static constexpr int BPAD = KLD;
  // CTA size follows the work per pass: small periods run many small CTAs per SM (measured: 64 threads up to
  // P = 48, 128 up to 80; e.g. P = 40: 74 -> 113 M conv/s, P = 80: 23.3 -> 28.5 M conv/s against 256 threads)
  static constexpr int THREADS = P <= 48 ? 64 : P <= 80 ? 128 : 256;
  static constexpr int A_LINES = THREADS / R2;
  static constexpr int A_ROUNDS = (NK + A_LINES - 1) / A_LINES;
  static constexpr int N0 = R2 / 2 + 1;          // half-spectrum columns in the p == 0 family
  static constexpr bool TW_IN_REGS = R1 <= 12;   // R1-1 complex twiddles per thread
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
};

template <int R1, int R2>
__global__ void __launch_bounds__((Geo<R1, R2>::THREADS), (Geo<R1, R2>::THREADS == 64 ? 8 : Geo<R1, R2>::THREADS == 128 ? 4 : Geo<R1, R2>::smem_bytes <= 110 * 1024 ? 2 : 1))
conv_kernel(const __grid_constant__ Params p) {
  typedef Geo<R1, R2> G;
  constexpr int HALF = G::HALF, H1 = G::H1, NK = G::NK, LD = G::LD, KLD = G::KLD, BPAD = G::BPAD;
  constexpr int THREADS = G::THREADS, A_LINES = G::A_LINES, A_ROUNDS = G::A_ROUNDS;
  cplx* Z = reinterpret_cast<cplx*>(b2c_smem_t2);
  const int t = threadIdx.x;
  const bool act = t < R2 * A_LINES;
  const int j = t / A_LINES, la = t - j * A_LINES;   // radix-R1 passes: lanes walk adjacent lines
  const int jj = act ? j : 0;
  cplx twr[G::TW_IN_REGS ? R1 : 1];
  if (G::TW_IN_REGS) {
#pragma unroll
    for (int q = 1; q < R1; ++q) twr[q % (G::TW_IN_REGS ? R1 : 1)] = __ldg(&p.tw[jj * q]);
  }
#define B2C_TW(q) (G::TW_IN_REGS ? twr[(q) % (G::TW_IN_REGS ? R1 : 1)] : __ldg(&p.tw[jj * (q)]))

  const size_t src_sz = (size_t)p.hs * p.ws, dst_sz = (size_t)p.hd * p.wd;
  for (long img = blockIdx.x; img < p.n_images; img += gridDim.x) {
    const double* __restrict__ x = p.src + (size_t)img * src_sz;
    double* __restrict__ y = p.dst + (size_t)img * dst_sz;

    if (t == 0 && img + gridDim.x < p.n_images) {   // next image -> L2
      const double* nx = p.src + (size_t)(img + gridDim.x) * src_sz;
      const unsigned bytes = (unsigned)((src_sz * sizeof(double)) & ~(size_t)15);
      if ((reinterpret_cast<size_t>(nx) & 15) == 0 && bytes >= 16)
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(nx), "r"(bytes) : "memory");
    }

    // ---- pass 1: vertical forward radix-R1, global -> Z
#pragma unroll 1
    for (int rnd = 0; rnd < A_ROUNDS; ++rnd) {
      const int line = rnd * A_LINES + la;
      if (act && line < HALF) {
        const bool c0 = line < p.ws_in, c1 = line + HALF < p.ws_in;
        cplx v[R1];
        if ((p.sh_h | p.sh_w) == 0) {
          const double* xr = x + j * p.ws + line;
          const int step = R2 * p.ws;
#pragma unroll
          for (int q = 0; q < R1; ++q) {
            const bool in = j + R2 * q < p.hs;
            v[q].x = (in && c0) ? __ldg(xr + q * step) : 0.0;
            v[q].y = (in && c1) ? __ldg(xr + q * step + HALF) : 0.0;
          }
        } else {   // source extended periodically by (sh_h, sh_w) samples in front
          int sc0 = line - p.sh_w, sc1 = line + HALF - p.sh_w;
          if (sc0 < 0) sc0 += p.ws;
          if (sc1 < 0) sc1 += p.ws;
#pragma unroll
          for (int q = 0; q < R1; ++q) {
            const int r = j + R2 * q;
            int sr = r - p.sh_h;
            if (sr < 0) sr += p.hs;
            const bool in = r < p.hs_in;
            v[q].x = (in && c0) ? __ldg(x + sr * p.ws + sc0) : 0.0;
            v[q].y = (in && c1) ? __ldg(x + sr * p.ws + sc1) : 0.0;
          }
        }
        Big<R1>::run(v);
        cplx* z = Z + line * LD + j;
        z[0] = v[0];
#pragma unroll
        for (int q = 1; q < R1; ++q) z[R2 * q] = cmul(v[q], B2C_TW(q));
      }
    }
    __syncthreads();

    // ---- pass 2: vertical forward radix-R2 (task = block pb of line `line`)
#pragma unroll 1
    for (int task = t; task < R1 * BPAD; task += THREADS) {
      const int pb = task / BPAD, line = task - pb * BPAD;
      if (line >= HALF) continue;
      cplx* z = Z + line * LD + R2 * pb;
      cplx v[R2];
#pragma unroll
      for (int s = 0; s < R2; ++s) v[s] = z[s];
      Big<R2>::run(v);
#pragma unroll
      for (int s = 0; s < R2; ++s) z[s] = v[s];
    }
    __syncthreads();

    // ---- pass 3: horizontal, Hermitian unpack + forward radix-R1
#pragma unroll 1
    for (int rnd = 0; rnd < A_ROUNDS; ++rnd) {
      const int kk = rnd * A_LINES + la;
      if (act && kk < NK) {
        const int rep = G::rep_of(kk), par = G::par_of(kk);
        const bool self = G::self_of(kk);
        cplx* zr = Z + j * LD + rep;
        cplx* zp = Z + j * LD + par;
        cplx v[R1];
#pragma unroll
        for (int q = 0; q < H1; ++q) {
          const cplx z1 = zr[R2 * q * LD];
          const cplx z2 = self ? z1 : zp[R2 * q * LD];
          v[q] = make_double2(z1.x + z2.x, z1.y - z2.y);          // 2 * FFT(column c)       at this frequency
          v[q + H1] = make_double2(z1.y + z2.y, z2.x - z1.x);     // 2 * FFT(column c + P/2)
        }
        Big<R1>::run(v);
        zr[0] = v[0];
#pragma unroll
        for (int q = 1; q < H1; ++q) zr[R2 * q * LD] = cmul(v[q], B2C_TW(q));
#pragma unroll
        for (int q = H1; q < R1; ++q) zp[R2 * (q - H1) * LD] = cmul(v[q], B2C_TW(q));
      }
    }
    __syncthreads();

    // ---- pass 4: horizontal radix-R2 forward, spectrum product, radix-R2 backward
#pragma unroll 1
    for (int task = t; task < R1 * BPAD; task += THREADS) {
      const int pb = task / BPAD, kk = task - pb * BPAD;
      if (kk >= NK) continue;
      const int col = pb < H1 ? G::rep_of(kk) : G::par_of(kk);
      cplx* z = Z + (R2 * (pb < H1 ? pb : pb - H1)) * LD + col;
      const cplx* __restrict__ ks = p.kspec + (R2 * pb) * KLD + kk;
      cplx v[R2];
#pragma unroll
      for (int s = 0; s < R2; ++s) v[s] = z[s * LD];
      Big<R2>::run(v);
#pragma unroll
      for (int s = 0; s < R2; ++s) v[s] = cswap(cmul(v[s], __ldg(ks + s * KLD)));   // product, swapped for the inverse
      Big<R2>::run(v);
#pragma unroll
      for (int s = 0; s < R2; ++s) z[s * LD] = cswap(v[s]);
    }
    __syncthreads();

    // ---- pass 5: horizontal backward radix-R1, repack two real columns into one complex line
#pragma unroll 1
    for (int rnd = 0; rnd < A_ROUNDS; ++rnd) {
      const int kk = rnd * A_LINES + la;
      if (act && kk < NK) {
        const int rep = G::rep_of(kk), par = G::par_of(kk);
        const bool self = G::self_of(kk);
        cplx* zr = Z + j * LD + rep;
        cplx* zp = Z + j * LD + par;
        cplx v[R1];
        v[0] = cswap(zr[0]);
#pragma unroll
        for (int q = 1; q < H1; ++q) v[q] = cswap(cmulc(zr[R2 * q * LD], B2C_TW(q)));
#pragma unroll
        for (int q = H1; q < R1; ++q) v[q] = cswap(cmulc(zp[R2 * (q - H1) * LD], B2C_TW(q)));
        Big<R1>::run(v);
#pragma unroll
        for (int q = 0; q < H1; ++q) {
          const cplx y1 = cswap(v[q]), y2 = cswap(v[q + H1]);
          zr[R2 * q * LD] = make_double2(y1.x - y2.y, y1.y + y2.x);              // Y1 + i*Y2 at +freq
          if (!self) zp[R2 * q * LD] = make_double2(y1.x + y2.y, y2.x - y1.y);   // conj(Y1) + i*conj(Y2) at -freq
        }
      }
    }
    __syncthreads();

    // ---- pass 6: vertical backward radix-R2
#pragma unroll 1
    for (int task = t; task < R1 * BPAD; task += THREADS) {
      const int pb = task / BPAD, line = task - pb * BPAD;
      if (line >= HALF) continue;
      cplx* z = Z + line * LD + R2 * pb;
      cplx v[R2];
#pragma unroll
      for (int s = 0; s < R2; ++s) v[s] = cswap(z[s]);
      Big<R2>::run(v);
#pragma unroll
      for (int s = 0; s < R2; ++s) z[s] = cswap(v[s]);
    }
    __syncthreads();

    // ---- pass 7: vertical backward radix-R1, crop, Z -> global
#pragma unroll 1
    for (int rnd = 0; rnd < A_ROUNDS; ++rnd) {
      const int line = rnd * A_LINES + la;
      if (act && line < HALF) {
        const cplx* z = Z + line * LD + j;
        cplx v[R1];
        v[0] = cswap(z[0]);
#pragma unroll
        for (int q = 1; q < R1; ++q) v[q] = cswap(cmulc(z[R2 * q], B2C_TW(q)));
        Big<R1>::run(v);   // results: re/im swapped => v[q].y is column `line`, v[q].x is column `line + P/2`
        const int j0 = line - p.off_w, j1 = line + HALF - p.off_w;
        const bool o0 = j0 >= 0 && j0 < p.wd, o1 = j1 >= 0 && j1 < p.wd;
        const int i0 = j - p.off_h;
        double* yr = y + i0 * p.wd + j0;
        const int step = R2 * p.wd;
#pragma unroll
        for (int q = 0; q < R1; ++q) {
          const bool in = (unsigned)(i0 + R2 * q) < (unsigned)p.hd;
          if (in && o0) yr[q * step] = v[q].y;
          if (in && o1) yr[q * step + HALF] = v[q].x;
        }
      }
    }
    // no barrier: the next image's pass 1 writes exactly the slots this thread just read
  }
#undef B2C_TW
}

// Kernel spectrum K[pos][kk] = 0.5/(P*P) * DFT2(kernel)[u(kk)][v(pos)], straight from the taps.
template <int R1, int R2>
__global__ void __launch_bounds__(512, 1) kspec_kernel(const double* __restrict__ ker, int hk, int wk,
                                                       const cplx* __restrict__ tw, cplx* __restrict__ out) {
  typedef Geo<R1, R2> G;
  constexpr int P = G::P, NK = G::NK, KLD = G::KLD;
  cplx* T1 = reinterpret_cast<cplx*>(b2c_smem_t2);  // [wke][NK]
  const int hke = min(hk, P), wke = min(wk, P);
  for (int idx = threadIdx.x; idx < wke * NK; idx += blockDim.x) {
    const int b = idx / NK, kk = idx - b * NK;
    const int u = G::freq_of_pos(G::rep_of(kk));
    cplx acc = make_double2(0.0, 0.0);
    for (int a = 0; a < hke; ++a) {
      double tap = 0.0;   // taps congruent modulo the period add up (convolution_fftw.h:185-187)
      for (int aa = a; aa < hk; aa += P)
        for (int bb = b; bb < wk; bb += P) tap += ker[(size_t)aa * wk + bb];
      const cplx w = tw[(u * a) % P];
      acc.x = fma(tap, w.x, acc.x);
      acc.y = fma(tap, w.y, acc.y);
    }
    T1[idx] = acc;
  }
  __syncthreads();
  const double scale = 0.5 / ((double)P * (double)P);
  for (int idx = threadIdx.x; idx < P * NK; idx += blockDim.x) {
    const int pos = idx / NK, kk = idx - pos * NK;
    const int v = G::freq_of_pos(pos);
    cplx acc = make_double2(0.0, 0.0);
    for (int b = 0; b < wke; ++b) {
      const cplx a = T1[b * NK + kk];
      const cplx w = tw[(v * b) % P];
      acc.x += a.x * w.x - a.y * w.y;
      acc.y += a.x * w.y + a.y * w.x;
    }
    out[pos * KLD + kk] = make_double2(acc.x * scale, acc.y * scale);
  }
}

template <int R1, int R2>
cudaError_t run(const Params& p, const double* d_ker, int hk, int wk, int sms, cudaStream_t st, int* launches) {
  typedef Geo<R1, R2> G;
  const size_t ksmem = sizeof(cplx) * (size_t)(wk < G::P ? wk : G::P) * G::NK;
  cudaFuncSetAttribute(kspec_kernel<R1, R2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ksmem);
  if (d_ker) {   // nullptr: the spectrum of the previous call is still valid (single-call kernel cache)
    kspec_kernel<R1, R2><<<1, 512, ksmem, st>>>(d_ker, hk, wk, p.tw, const_cast<cplx*>(p.kspec));
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    ++*launches;
  }
  int occ = 0;
  cudaFuncSetAttribute(conv_kernel<R1, R2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G::smem_bytes);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, conv_kernel<R1, R2>, G::THREADS, G::smem_bytes);
  if (occ < 1) return cudaErrorLaunchOutOfResources;
  long grid = (long)sms * occ;
  if (grid > p.n_images) grid = p.n_images;
  conv_kernel<R1, R2><<<(int)grid, G::THREADS, G::smem_bytes, st>>>(p);
  ++*launches;
  return cudaGetLastError();
}

}  // namespace t2

// ---- the menu: square periods with a tuned kernel --------------------------------------------------
#define B2C_T2_MENU(X)                                                                                         \
  X(8, 2) X(10, 2) X(8, 3) X(4, 7) X(8, 4) X(6, 6) X(8, 5) X(6, 7) X(8, 6) X(10, 5) X(6, 9) X(8, 7) X(10, 6) X(8, 8) \
  X(10, 7) X(8, 9) X(10, 8) X(12, 7) X(10, 9) X(12, 8) X(10, 10) X(12, 9) X(14, 8) X(12, 10) X(14, 9) X(16, 8)

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
                          const cplx* tw, int sms, cudaStream_t st, int* launches) {
  t2::Params p;
  p.hs = hs; p.ws = ws; p.hd = hd; p.wd = wd; p.off_h = off_h; p.off_w = off_w;
  p.hs_in = hs_in; p.ws_in = ws_in; p.sh_h = sh_h; p.sh_w = sh_w;
  p.src = d_src; p.dst = d_dst; p.n_images = n; p.kspec = kspec; p.tw = tw;
  switch (P) {
#define X(a, b) case (a) * (b): return t2::run<a, b>(p, d_ker, hk, wk, sms, st, launches);
    B2C_T2_MENU(X)
#undef X
  }
  return cudaErrorInvalidValue;
}

}  // namespace b2c
