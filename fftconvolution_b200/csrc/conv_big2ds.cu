// Strip-mode fused FFT convolution with a COMPILE-TIME radix schedule: the tuned form of
// conv_big2d.cu for the period of BASELINE config 3 (100x100 (*) 100x100 CIRCULAR_FULL -> 200 x 200,
// per-image kernels).  Replaces fftw_circular_convolution + crop (reference convolution_fftw.h:169-284).
//
// Same three phases and the same L2-resident slab as conv_big2d.cu (rows -> half spectra -> column
// strips with the product -> rows), but every radix pass is instantiated for its literal length,
// block size and radix, and the work is laid out so that
//   * the 32 lanes of a warp walk adjacent LINES (rows of a chunk / columns of a strip) and a warp
//     owns one butterfly index: block, offset and the R-1 twiddles are warp-uniform, the twiddles are
//     constant-bank operands of the FP64 instructions (no registers, no loads, no index arithmetic);
//   * the copy loops run one warp per row with lanes along the contiguous dimension, four
//     independent loads in flight, and no integer division.
#include "conv_params.h"
#include "tw200_table.h"

namespace b2c {

extern __shared__ __align__(16) unsigned char b2c_smem_bigs[];

namespace bigs {

#ifndef B2C_BIGS_TRACE
#define B2C_BIGS_TRACE 0
#endif
#if B2C_BIGS_TRACE
__device__ long long g_acc[16];   // debug variant: cycles accumulated per phase by CTA 0 / thread 0
#define B2C_T(slot) do { if (blockIdx.x == 0 && threadIdx.x == 0) { long long c_; asm volatile("mov.u64 %0, %%clock64;" : "=l"(c_)); g_acc[slot] += c_ - tprev; tprev = c_; } } while (0)
#else
#define B2C_T(slot)
#endif

template <int P> struct TwTable;
template <> struct TwTable<200> { static __device__ __forceinline__ cplx at(int k) { return c_tw200[k]; } };

// One radix-R stage of a length-P transform on `nlines` lines (element k of line l at W[l*ls + k*es]),
// block length N, lanes = lines, warps = butterfly indices.
template <int P, int R, int N, bool INV>
__device__ __forceinline__ void stage(cplx* W, int es, int ls, int nlines, int warp, int nwarps, int lane) {
  constexpr int M = N / R, PER_LINE = P / R, TSTEP = P / N;
  for (int t = warp; t < PER_LINE; t += nwarps) {      // warp-uniform
    const int b = t / M, j = t - b * M;
    cplx* base = W + (b * N + j) * es;
    for (int line = lane; line < nlines; line += 32) {
      cplx* p = base + line * ls;
      cplx v[R];
#pragma unroll
      for (int q = 0; q < R; ++q) v[q] = p[q * M * es];
      if (INV) {
        if (M > 1) {
#pragma unroll
          for (int q = 1; q < R; ++q) v[q] = cmulc(v[q], TwTable<P>::at(j * q * TSTEP));
        }
#pragma unroll
        for (int q = 0; q < R; ++q) v[q] = cswap(v[q]);
        Bfly<R>::run(v);
#pragma unroll
        for (int q = 0; q < R; ++q) v[q] = cswap(v[q]);
      } else {
        Bfly<R>::run(v);
        if (M > 1) {
#pragma unroll
          for (int q = 1; q < R; ++q) v[q] = cmul(v[q], TwTable<P>::at(j * q * TSTEP));
        }
      }
#pragma unroll
      for (int q = 0; q < R; ++q) p[q * M * es] = v[q];
    }
  }
  __syncthreads();
}

template <int P, int R0, int R1, int R2>
struct Fft {
  static __device__ __forceinline__ void forward(cplx* W, int es, int ls, int nlines, int warp, int nwarps, int lane) {
    stage<P, R0, P, false>(W, es, ls, nlines, warp, nwarps, lane);
    stage<P, R1, P / R0, false>(W, es, ls, nlines, warp, nwarps, lane);
    stage<P, R2, P / R0 / R1, false>(W, es, ls, nlines, warp, nwarps, lane);
  }
  static __device__ __forceinline__ void backward(cplx* W, int es, int ls, int nlines, int warp, int nwarps, int lane) {
    stage<P, R2, P / R0 / R1, true>(W, es, ls, nlines, warp, nwarps, lane);
    stage<P, R1, P / R0, true>(W, es, ls, nlines, warp, nwarps, lane);
    stage<P, R0, P, true>(W, es, ls, nlines, warp, nwarps, lane);
  }
};

__device__ __forceinline__ double wrapped_tap(const BigParams& p, const double* ker, int r, int c) {
  double acc = 0.0;   // taps congruent modulo the period add up (convolution_fftw.h:185-187)
  for (int i = r; i < p.hk; i += p.kwrap_h)
    for (int j = c; j < p.wk; j += p.kwrap_w) acc += ker[(size_t)i * p.wk + j];
  return acc;
}

template <int P, int R0, int R1, int R2>
__global__ void __launch_bounds__(512, 1) conv_kernel(const __grid_constant__ BigParams p) {
  typedef Fft<P, R0, R1, R2> F;
  cplx* W = reinterpret_cast<cplx*>(b2c_smem_bigs);
  const int lane = threadIdx.x & 31;
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);   // uniform for the compiler
  const int nwarps = blockDim.x >> 5;
  const int hke = min(p.hk, p.kwrap_h), wke = min(p.wk, p.kwrap_w);
  const int rows_in = max(p.hs, hke);
  const int HW = p.half_w, ldw = p.ldw, ldc = p.ldc;
  const bool plain_kernel = (p.hk <= p.kwrap_h) && (p.wk <= p.kwrap_w);
  const size_t src_sz = (size_t)p.hs * p.ws, dst_sz = (size_t)p.hd * p.wd;
  cplx* slabS = p.slab + (size_t)blockIdx.x * (size_t)(p.hs + hke + p.rows_y) * HW;
  cplx* slabK = slabS + (size_t)p.hs * HW;
  cplx* slabY = slabK + (size_t)hke * HW;
  const int* __restrict__ posf = p.aw.posf;

#if B2C_BIGS_TRACE
  long long tprev; asm volatile("mov.u64 %0, %%clock64;" : "=l"(tprev));
#endif
  for (long img = blockIdx.x; img < p.n_images; img += gridDim.x) {
    const double* a = p.src + (size_t)img * src_sz;
    const double* k = p.ker + (size_t)img * p.ker_stride;

    // ---------------- phase R: z_r = s_r + i*k_r, forward rows, Hermitian split -> slab
    for (int r0 = 0; r0 < rows_in; r0 += p.rows_chunk) {
      const int nr = min(p.rows_chunk, rows_in - r0);
      for (int rr = warp; rr < nr; rr += nwarps) {
        const int r = r0 + rr;
        cplx* wrow = W + rr * ldw;
        const double* arow = a + (size_t)r * p.ws;
        const double* krow = k + (size_t)r * p.wk;
        constexpr int NCH = (P + 31) / 32;       // all loads of a row are in flight before the first store
        cplx v[NCH];
#pragma unroll
        for (int u = 0; u < NCH; ++u) {
          const int c = lane + 32 * u;
          v[u] = make_double2(0.0, 0.0);
          if (r < p.hs && c < p.ws) v[u].x = __ldg(arow + c);
          if (r < hke && c < wke) v[u].y = plain_kernel ? __ldg(krow + c) : wrapped_tap(p, k, r, c);
        }
#pragma unroll
        for (int u = 0; u < NCH; ++u)
          if (lane + 32 * u < P) wrow[lane + 32 * u] = v[u];
      }
      __syncthreads();
      B2C_T(0);
      F::forward(W, 1, ldw, nr, warp, nwarps, lane);
      B2C_T(1);
      // S(v) = (Z(v) + conj Z(-v)) / 2,  K(v) = -i (Z(v) - conj Z(-v)) / 2
      for (int rr = warp; rr < nr; rr += nwarps) {
        const int r = r0 + rr;
        const cplx* wrow = W + rr * ldw;
        for (int v = lane; v < HW; v += 32) {
          const cplx z1 = wrow[__ldg(posf + v)];
          const cplx z2 = wrow[__ldg(posf + (v ? P - v : 0))];
          if (r < p.hs) slabS[(size_t)r * HW + v] = make_double2(0.5 * (z1.x + z2.x), 0.5 * (z1.y - z2.y));
          if (r < hke) slabK[(size_t)r * HW + v] = make_double2(0.5 * (z1.y + z2.y), 0.5 * (z2.x - z1.x));
        }
      }
      __syncthreads();
      B2C_T(2);
    }

    // ---------------- phase C: strips of half-spectrum columns: both column transforms, product, backward
    for (int v0 = 0; v0 < HW; v0 += p.cols_strip) {
      const int nc = min(p.cols_strip, HW - v0);
      for (int rb = 4 * warp; rb < P; rb += 4 * nwarps) {     // four rows per warp trip: 8+ loads in flight per lane
        for (int c0 = lane; c0 < 2 * nc; c0 += 64) {
          cplx v[4][2];
#pragma unroll
          for (int rr = 0; rr < 4; ++rr) {
            const int r = rb + rr;
#pragma unroll
            for (int u = 0; u < 2; ++u) {
              const int c = c0 + 32 * u;
              v[rr][u] = make_double2(0.0, 0.0);
              if (c < nc) { if (r < p.hs) v[rr][u] = slabS[(size_t)r * HW + v0 + c]; }
              else if (c < 2 * nc) { if (r < hke) v[rr][u] = slabK[(size_t)r * HW + v0 + c - nc]; }
            }
          }
#pragma unroll
          for (int rr = 0; rr < 4; ++rr)
#pragma unroll
            for (int u = 0; u < 2; ++u)
              if (rb + rr < P && c0 + 32 * u < 2 * nc) W[(rb + rr) * ldc + c0 + 32 * u] = v[rr][u];
        }
      }
      __syncthreads();
      B2C_T(3);
      F::forward(W, ldc, 1, 2 * nc, warp, nwarps, lane);
      B2C_T(4);
      for (int r = warp; r < P; r += nwarps) {   // product, 1/N folded in (convolution_fftw.h:195-204, 210-211)
        cplx* wrow = W + r * ldc;
        for (int c = lane; c < nc; c += 32) {
          const cplx y = cmul(wrow[c], wrow[nc + c]);
          wrow[c] = make_double2(y.x * p.scale, y.y * p.scale);
        }
      }
      __syncthreads();
      B2C_T(5);
      F::backward(W, ldc, 1, nc, warp, nwarps, lane);
      B2C_T(6);
      for (int r = warp; r < p.rows_y; r += nwarps) {
        const cplx* wrow = W + r * ldc;
        for (int c = lane; c < nc; c += 32) slabY[(size_t)r * HW + v0 + c] = wrow[c];
      }
      __syncthreads();
      B2C_T(7);
    }

    // ---------------- phase R': W = Y_r + i*Y_{r+1} (Hermitian-extended), backward rows, crop -> dst
    double* d = p.dst + (size_t)img * dst_sz;
    const int pairs = (p.hd + 1) / 2;
    for (int q0 = 0; q0 < pairs; q0 += p.rows_chunk) {
      const int nq = min(p.rows_chunk, pairs - q0);
      for (int qq = warp; qq < nq; qq += nwarps) {
        const int i0 = 2 * (q0 + qq), i1 = i0 + 1;
        const cplx* y0row = slabY + (size_t)(p.off_h + i0) * HW;
        const cplx* y1row = slabY + (size_t)(p.off_h + i1) * HW;
        const bool has1 = i1 < p.hd;
        cplx* wrow = W + qq * ldw;
        constexpr int NCH = (P + 31) / 32;
        cplx y0[NCH], y1[NCH];
        int pos[NCH];
#pragma unroll
        for (int u = 0; u < NCH; ++u) {
          const int v = lane + 32 * u;
          pos[u] = -1;
          if (v < P) {
            const int vv = v < HW ? v : P - v;
            y0[u] = y0row[vv];
            y1[u] = has1 ? y1row[vv] : make_double2(0.0, 0.0);
            if (v >= HW) { y0[u].y = -y0[u].y; y1[u].y = -y1[u].y; }   // Hermitian mirror
            pos[u] = __ldg(posf + v);
          }
        }
#pragma unroll
        for (int u = 0; u < NCH; ++u)
          if (pos[u] >= 0) wrow[pos[u]] = make_double2(y0[u].x - y1[u].y, y0[u].y + y1[u].x);
      }
      __syncthreads();
      B2C_T(8);
      F::backward(W, 1, ldw, nq, warp, nwarps, lane);
      B2C_T(9);
      for (int ri = warp; ri < 2 * nq; ri += nwarps) {
        const int i = 2 * q0 + ri;
        if (i < p.hd) {
          const cplx* wrow = W + (ri >> 1) * ldw + p.off_w;
          double* drow = d + (size_t)i * p.wd;
          if (ri & 1) { for (int j = lane; j < p.wd; j += 32) drow[j] = wrow[j].y; }
          else        { for (int j = lane; j < p.wd; j += 32) drow[j] = wrow[j].x; }
        }
      }
      __syncthreads();
      B2C_T(10);
    }
  }
}

}  // namespace bigs

// true when a compile-time strip kernel exists for this (square) period and radix schedule
#if B2C_BIGS_TRACE
extern "C" int b2c_debug_bigs_read(long long* host) { return (int)cudaMemcpyFromSymbol(host, bigs::g_acc, sizeof(long long) * 16); }
#endif

bool big2ds_has(int Ph, int Pw, const int* rad_h, int ns_h, const int* rad_w, int ns_w) {
  auto is200 = [](int P, const int* r, int ns) { return P == 200 && ns == 3 && r[0] == 5 && r[1] == 5 && r[2] == 8; };
  return is200(Ph, rad_h, ns_h) && is200(Pw, rad_w, ns_w);
}

cudaError_t launch_conv_big2ds(const BigParams& p, int grid, size_t smem, cudaStream_t st) {
  auto kern = bigs::conv_kernel<200, 5, 5, 8>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  kern<<<grid, 512, smem, st>>>(p);
  return cudaGetLastError();
}

}  // namespace b2c
