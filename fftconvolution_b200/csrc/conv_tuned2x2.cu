// Experiment: the seven passes of conv_tuned2.cu with TWO images per CTA processed in lockstep --
// one CTA of 256 threads per SM, two work buffers, every thread runs each of its tasks on both images with the two
// instruction streams interleaved (twice the instruction-level parallelism, shared twiddles / addresses / kernel-
// spectrum loads / barriers, twice the loads in flight per latency).  Shared kernel, no periodic extension.
// Replaces fftw_circular_convolution + crop (reference convolution_fftw.h:169-284) like the kernel it derives from.
#include "conv_params.h"

namespace b2c {

extern __shared__ __align__(16) unsigned char b2c_smem_t2x2[];

namespace t2 {
// conv_tuned2.cu
struct Params {
  int hs, ws, hd, wd, off_h, off_w;
  int hs_in, ws_in, sh_h, sh_w;
  const double* src;
  double* dst;
  long n_images;
  const cplx* kspec;
  const cplx* tw;
  const double* ker;
  int hk, wk;
  cplx* kscratch;
};
}  // namespace t2

namespace t2x2 {

using t2::Params;
template <int R> using Big = BigBfly<R>;
constexpr int NI = 2;

template <int R1, int R2>
struct Geo {
  static constexpr int P = R1 * R2, HALF = P / 2, H1 = R1 / 2, NK = P / 2 + 1;
  static constexpr int LD = ((P + 2) | 1) + ((((P + 2) | 1) % 8 == 1) ? 2 : 0);
  static constexpr int KLD = (NK + 7) / 8 * 8, BPAD = KLD;
  static constexpr int THREADS = 256, A_LINES = THREADS / R2, A_ROUNDS = (NK + A_LINES - 1) / A_LINES, N0 = R2 / 2 + 1;
  static __host__ __device__ constexpr int rep_of(int kk) { return kk < N0 ? kk : R2 + (kk - N0); }
  static __host__ __device__ constexpr bool self_of(int kk) {
    return kk == 0 || (kk < N0 ? (R2 % 2 == 0 && kk == R2 / 2) : (P + R2 - 1 - rep_of(kk) == rep_of(kk)));
  }
  static __host__ __device__ constexpr int par_of(int kk) {
    return kk == 0 ? P : (self_of(kk) ? P + 1 : (kk < N0 ? R2 - kk : P + R2 - 1 - rep_of(kk)));
  }
  static constexpr int ZSZ = HALF * LD;
  static constexpr size_t smem_bytes = sizeof(cplx) * (size_t)NI * ZSZ;
};

template <int R1, int R2>
__global__ void __launch_bounds__(256, 1) conv_kernel(const __grid_constant__ Params p) {
  typedef Geo<R1, R2> G;
  constexpr int HALF = G::HALF, H1 = G::H1, NK = G::NK, LD = G::LD, KLD = G::KLD, BPAD = G::BPAD, ZSZ = G::ZSZ;
  constexpr int THREADS = G::THREADS, A_LINES = G::A_LINES, A_ROUNDS = G::A_ROUNDS;
  cplx* Z = reinterpret_cast<cplx*>(b2c_smem_t2x2);
  const int t = threadIdx.x;
  const bool act = t < R2 * A_LINES;
  const int j = t / A_LINES, la = t - j * A_LINES;
  const int jj = act ? j : 0;
  cplx twr[R1];
#pragma unroll
  for (int q = 1; q < R1; ++q) twr[q] = __ldg(&p.tw[jj * q]);

  const size_t src_sz = (size_t)p.hs * p.ws, dst_sz = (size_t)p.hd * p.wd;
  const long pairs = (p.n_images + 1) / 2;
  for (long pr = blockIdx.x; pr < pairs; pr += gridDim.x) {
    const long img = 2 * pr;
    const bool two = img + 1 < p.n_images;
    const double* __restrict__ x0 = p.src + (size_t)img * src_sz;
    double* __restrict__ y0 = p.dst + (size_t)img * dst_sz;
    if (t == 0 && pr + gridDim.x < pairs) {
      const double* nx = p.src + (size_t)(2 * (pr + gridDim.x)) * src_sz;
      const size_t nimg = 2 * (pr + gridDim.x) + 1 < p.n_images ? 2 : 1;
      l2_prefetch(nx, nimg * src_sz * sizeof(double));
    }

    // ---- pass 1
#pragma unroll 1
    for (int rnd = 0; rnd < A_ROUNDS; ++rnd) {
      const int line = rnd * A_LINES + la;
      if (act && line < HALF) {
        const bool c0 = line < p.ws, c1 = line + HALF < p.ws;
        cplx v[NI][R1];
        const int step = R2 * p.ws;
#pragma unroll
        for (int im = 0; im < NI; ++im) {
          const double* xr = x0 + im * src_sz + j * p.ws + line;
          const bool on = im == 0 || two;
#pragma unroll
          for (int q = 0; q < R1; ++q) {
            const bool in = on && j + R2 * q < p.hs;
            v[im][q].x = (in && c0) ? __ldg(xr + q * step) : 0.0;
            v[im][q].y = (in && c1) ? __ldg(xr + q * step + HALF) : 0.0;
          }
        }
#pragma unroll
        for (int im = 0; im < NI; ++im) Big<R1>::run(v[im]);
#pragma unroll
        for (int im = 0; im < NI; ++im) {
          cplx* z = Z + im * ZSZ + line * LD + j;
          z[0] = v[im][0];
#pragma unroll
          for (int q = 1; q < R1; ++q) z[R2 * q] = cmul(v[im][q], twr[q]);
        }
      }
    }
    __syncthreads();

    // ---- pass 2
#pragma unroll 1
    for (int task = t; task < R1 * BPAD; task += THREADS) {
      const int pb = task / BPAD, line = task - pb * BPAD;
      if (line >= HALF) continue;
      cplx* z = Z + line * LD + R2 * pb;
      cplx v[NI][R2];
#pragma unroll
      for (int im = 0; im < NI; ++im)
#pragma unroll
        for (int s = 0; s < R2; ++s) v[im][s] = z[im * ZSZ + s];
#pragma unroll
      for (int im = 0; im < NI; ++im) Big<R2>::run(v[im]);
#pragma unroll
      for (int im = 0; im < NI; ++im)
#pragma unroll
        for (int s = 0; s < R2; ++s) z[im * ZSZ + s] = v[im][s];
    }
    __syncthreads();

    // ---- pass 3
#pragma unroll 1
    for (int rnd = 0; rnd < A_ROUNDS; ++rnd) {
      const int kk = rnd * A_LINES + la;
      if (act && kk < NK) {
        const int rep = G::rep_of(kk), par = G::par_of(kk);
        const bool self = G::self_of(kk);
        cplx* zr = Z + j * LD + rep;
        cplx* zp = Z + j * LD + par;
        cplx v[NI][R1];
#pragma unroll
        for (int im = 0; im < NI; ++im)
#pragma unroll
          for (int q = 0; q < H1; ++q) {
            const cplx z1 = zr[im * ZSZ + R2 * q * LD];
            const cplx z2 = self ? z1 : zp[im * ZSZ + R2 * q * LD];
            v[im][q] = make_double2(z1.x + z2.x, z1.y - z2.y);
            v[im][q + H1] = make_double2(z1.y + z2.y, z2.x - z1.x);
          }
#pragma unroll
        for (int im = 0; im < NI; ++im) Big<R1>::run(v[im]);
#pragma unroll
        for (int im = 0; im < NI; ++im) {
          zr[im * ZSZ] = v[im][0];
#pragma unroll
          for (int q = 1; q < H1; ++q) zr[im * ZSZ + R2 * q * LD] = cmul(v[im][q], twr[q]);
#pragma unroll
          for (int q = H1; q < R1; ++q) zp[im * ZSZ + R2 * (q - H1) * LD] = cmul(v[im][q], twr[q]);
        }
      }
    }
    __syncthreads();

    // ---- pass 4
#pragma unroll 1
    for (int task = t; task < R1 * BPAD; task += THREADS) {
      const int pb = task / BPAD, kk = task - pb * BPAD;
      if (kk >= NK) continue;
      const int col = pb < H1 ? G::rep_of(kk) : G::par_of(kk);
      cplx* z = Z + (R2 * (pb < H1 ? pb : pb - H1)) * LD + col;
      const cplx* __restrict__ ks = p.kspec + (R2 * pb) * KLD + kk;
      cplx v[NI][R2], w[R2];
#pragma unroll
      for (int s = 0; s < R2; ++s) w[s] = __ldg(ks + s * KLD);   // one spectrum load serves both images
#pragma unroll
      for (int im = 0; im < NI; ++im)
#pragma unroll
        for (int s = 0; s < R2; ++s) v[im][s] = z[im * ZSZ + s * LD];
#pragma unroll
      for (int im = 0; im < NI; ++im) Big<R2>::run(v[im]);
#pragma unroll
      for (int im = 0; im < NI; ++im)
#pragma unroll
        for (int s = 0; s < R2; ++s) v[im][s] = cswap(cmul(v[im][s], w[s]));
#pragma unroll
      for (int im = 0; im < NI; ++im) Big<R2>::run(v[im]);
#pragma unroll
      for (int im = 0; im < NI; ++im)
#pragma unroll
        for (int s = 0; s < R2; ++s) z[im * ZSZ + s * LD] = cswap(v[im][s]);
    }
    __syncthreads();

    // ---- pass 5
#pragma unroll 1
    for (int rnd = 0; rnd < A_ROUNDS; ++rnd) {
      const int kk = rnd * A_LINES + la;
      if (act && kk < NK) {
        const int rep = G::rep_of(kk), par = G::par_of(kk);
        const bool self = G::self_of(kk);
        cplx* zr = Z + j * LD + rep;
        cplx* zp = Z + j * LD + par;
        cplx v[NI][R1];
#pragma unroll
        for (int im = 0; im < NI; ++im) {
          v[im][0] = cswap(zr[im * ZSZ]);
#pragma unroll
          for (int q = 1; q < H1; ++q) v[im][q] = cswap(cmulc(zr[im * ZSZ + R2 * q * LD], twr[q]));
#pragma unroll
          for (int q = H1; q < R1; ++q) v[im][q] = cswap(cmulc(zp[im * ZSZ + R2 * (q - H1) * LD], twr[q]));
        }
#pragma unroll
        for (int im = 0; im < NI; ++im) Big<R1>::run(v[im]);
#pragma unroll
        for (int im = 0; im < NI; ++im)
#pragma unroll
          for (int q = 0; q < H1; ++q) {
            const cplx y1 = cswap(v[im][q]), y2 = cswap(v[im][q + H1]);
            zr[im * ZSZ + R2 * q * LD] = make_double2(y1.x - y2.y, y1.y + y2.x);
            if (!self) zp[im * ZSZ + R2 * q * LD] = make_double2(y1.x + y2.y, y2.x - y1.y);
          }
      }
    }
    __syncthreads();

    // ---- pass 6
#pragma unroll 1
    for (int task = t; task < R1 * BPAD; task += THREADS) {
      const int pb = task / BPAD, line = task - pb * BPAD;
      if (line >= HALF) continue;
      cplx* z = Z + line * LD + R2 * pb;
      cplx v[NI][R2];
#pragma unroll
      for (int im = 0; im < NI; ++im)
#pragma unroll
        for (int s = 0; s < R2; ++s) v[im][s] = cswap(z[im * ZSZ + s]);
#pragma unroll
      for (int im = 0; im < NI; ++im) Big<R2>::run(v[im]);
#pragma unroll
      for (int im = 0; im < NI; ++im)
#pragma unroll
        for (int s = 0; s < R2; ++s) z[im * ZSZ + s] = cswap(v[im][s]);
    }
    __syncthreads();

    // ---- pass 7
#pragma unroll 1
    for (int rnd = 0; rnd < A_ROUNDS; ++rnd) {
      const int line = rnd * A_LINES + la;
      if (act && line < HALF) {
        cplx v[NI][R1];
#pragma unroll
        for (int im = 0; im < NI; ++im) {
          const cplx* z = Z + im * ZSZ + line * LD + j;
          v[im][0] = cswap(z[0]);
#pragma unroll
          for (int q = 1; q < R1; ++q) v[im][q] = cswap(cmulc(z[R2 * q], twr[q]));
        }
#pragma unroll
        for (int im = 0; im < NI; ++im) Big<R1>::run(v[im]);
        const int j0 = line - p.off_w, j1 = line + HALF - p.off_w;
        const bool o0 = j0 >= 0 && j0 < p.wd, o1 = j1 >= 0 && j1 < p.wd;
        const int i0 = j - p.off_h;
        const int step = R2 * p.wd;
#pragma unroll
        for (int im = 0; im < NI; ++im) {
          if (im == 1 && !two) break;
          double* yr = y0 + im * dst_sz + i0 * p.wd + j0;
#pragma unroll
          for (int q = 0; q < R1; ++q) {
            const bool in = (unsigned)(i0 + R2 * q) < (unsigned)p.hd;
            if (in && o0) yr[q * step] = v[im][q].y;
            if (in && o1) yr[q * step + HALF] = v[im][q].x;
          }
        }
      }
    }
    // no barrier: the next pair's pass 1 writes exactly the slots this thread just read
  }
}

}  // namespace t2x2

// two images per CTA at period 108 (kernel spectrum already in `kspec`, layout of t2::kspec_kernel<12, 9>)
cudaError_t launch_tuned2x2_108(const t2::Params& p, int sms, cudaStream_t st) {
  typedef t2x2::Geo<12, 9> G;
  cudaFuncSetAttribute(t2x2::conv_kernel<12, 9>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G::smem_bytes);
  const long pairs = (p.n_images + 1) / 2;
  const int grid = (int)(pairs < sms ? pairs : sms);
  t2x2::conv_kernel<12, 9><<<grid, G::THREADS, G::smem_bytes, st>>>(p);
  return cudaGetLastError();
}

}  // namespace b2c
