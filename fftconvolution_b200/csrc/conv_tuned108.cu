// Tuned fused convolution kernel for period 108 x 108 (= 12 * 9 on both axes) -- the plan the
// headline workload gets (100x100 (*) 15x15 LINEAR_SAME needs a period >= 107).
// Replaces fftw_circular_convolution + crop (reference convolution_fftw.h:169-284) for that plan.
//
// One CTA owns one image; two CTAs are resident per SM (99.4 KB of shared memory each), so one
// loads/stores while the other computes.  The image is transformed VERTICALLY first so that the
// lanes of a warp walk adjacent image columns: global loads and stores are coalesced.
// Data flow, all in ONE shared buffer Z[54][115] complex ("line" = a pair of image columns):
//   1 vert. fwd radix-12 : global -> Z.  Columns c and c+54 ride one complex transform
//                          (z = x_c + i*x_{c+54}); zero padding happens in registers.
//   2 vert. fwd radix-9
//   3 horiz fwd radix-12 : Hermitian unpack of the two packed columns in registers, then butterfly.
//                          Only 55 of the 108 vertical frequencies exist (real input).
//   4 horiz radix-9 fwd, * cached kernel spectrum (1/N folded in), radix-9 bwd -- one pass.
//   5 horiz bwd radix-12 : butterfly, then repack columns c / c+54 into one complex line.
//   6 vert. bwd radix-9
//   7 vert. bwd radix-12 : Z -> global, cropping to the output window in registers.
// Every pass is in place: a task reads and writes the same slots, so a single __syncthreads()
// separates passes (6 per image).  In every pass adjacent lanes own adjacent LINES (stride
// LD = 115, odd) or adjacent columns of Z, so 128-bit shared accesses of a quarter warp fall into
// distinct bank groups.  The 11 twiddles a thread needs (w108^(j*p), j fixed per thread) live in
// registers.
#include "conv_params.h"

namespace b2c {

extern __shared__ __align__(16) unsigned char b2c_smem_t108[];

namespace t108 {

constexpr int P = 108;
constexpr int HALF = 54;
constexpr int LD = 115;          // odd stride (bank groups), chosen with tools/bank_sim_t108.py
constexpr int NK = 55;          // half-spectrum columns
constexpr int KLD = 56;         // kernel-spectrum row pitch (896 B = 7 cache lines)
constexpr int BPAD = 56;        // radix-9 passes index 56 lines per block so block changes fall on quarter-warp edges
constexpr int THREADS = 256;
constexpr int A_LINES = 28;     // radix-12 passes: 28 lines x 9 butterflies = 252 active threads per round

struct Params {
  int hs, ws, hd, wd, off_h, off_w;
  const double* src;
  double* dst;
  long n_images;
  const cplx* kspec;  // [P][KLD]: (scrambled horizontal position, half-spectrum vertical frequency kk)
  const cplx* tw;     // exp(-2 pi i k / 108)
};

// frequency held at scrambled position `pos` after the (12, 9) decimation-in-frequency passes
__host__ __device__ __forceinline__ int freq_of_pos(int pos) { return pos / 9 + 12 * (pos % 9); }
// half-spectrum column kk -> scrambled column position of its representative frequency
__host__ __device__ __forceinline__ int rep_of(int kk) { return kk <= 4 ? kk : kk + 4; }
// ... and the slot column that holds the Hermitian partner (108 / 109: spare columns of the two
// self-conjugate frequencies 0 and 54)
__host__ __device__ __forceinline__ int par_of(int kk) {
  return kk == 0 ? 108 : (kk <= 4 ? 9 - kk : (kk == 54 ? 109 : 112 - kk));
}

__device__ __forceinline__ void bfly9(cplx* v) {
  const cplx w1 = make_double2(0.76604444311897803520, -0.64278760968653932632);
  const cplx w2 = make_double2(0.17364817766693034885, -0.98480775301220805937);
  const cplx w4 = make_double2(-0.93969262078590838405, -0.34202014332566873304);
  cplx a0[3] = {v[0], v[3], v[6]};
  cplx a1[3] = {v[1], v[4], v[7]};
  cplx a2[3] = {v[2], v[5], v[8]};
  Bfly<3>::run(a0);
  Bfly<3>::run(a1);
  Bfly<3>::run(a2);
  a1[1] = cmul(a1[1], w1);
  a1[2] = cmul(a1[2], w2);
  a2[1] = cmul(a2[1], w2);
  a2[2] = cmul(a2[2], w4);
#pragma unroll
  for (int k1 = 0; k1 < 3; ++k1) {
    cplx s[3] = {a0[k1], a1[k1], a2[k1]};
    Bfly<3>::run(s);
    v[k1] = s[0];
    v[k1 + 3] = s[1];
    v[k1 + 6] = s[2];
  }
}

// 12 = 3 x 4 prime-factor (Good-Thomas) butterfly: no internal twiddles.
__device__ __forceinline__ void bfly12(cplx* v) {
  cplx u[3][4];
#pragma unroll
  for (int n1 = 0; n1 < 3; ++n1) {
    cplx t[4];
#pragma unroll
    for (int n2 = 0; n2 < 4; ++n2) t[n2] = v[(4 * n1 + 3 * n2) % 12];
    Bfly<4>::run(t);
#pragma unroll
    for (int k2 = 0; k2 < 4; ++k2) u[n1][k2] = t[k2];
  }
#pragma unroll
  for (int k2 = 0; k2 < 4; ++k2) {
    cplx s[3] = {u[0][k2], u[1][k2], u[2][k2]};
    Bfly<3>::run(s);
#pragma unroll
    for (int k1 = 0; k1 < 3; ++k1) v[(4 * k1 + 9 * k2) % 12] = s[k1];
  }
}

template <int R>
__device__ __forceinline__ void swap_all(cplx* v) {
#pragma unroll
  for (int q = 0; q < R; ++q) v[q] = cswap(v[q]);
}

__global__ void __launch_bounds__(THREADS, 2) conv_kernel(const __grid_constant__ Params p) {
  cplx* Z = reinterpret_cast<cplx*>(b2c_smem_t108);
  const int t = threadIdx.x;
  const bool act = t < 9 * A_LINES;
  const int j = t / A_LINES, la = t - j * A_LINES;   // radix-12 passes: lanes walk adjacent lines
  cplx tw[12];
#pragma unroll
  for (int q = 1; q < 12; ++q) tw[q] = __ldg(&p.tw[(act ? j : 0) * q]);

  const size_t src_sz = (size_t)p.hs * p.ws, dst_sz = (size_t)p.hd * p.wd;
  for (long img = blockIdx.x; img < p.n_images; img += gridDim.x) {
    const double* __restrict__ x = p.src + (size_t)img * src_sz;
    double* __restrict__ y = p.dst + (size_t)img * dst_sz;

    // next image of this CTA -> L2 while this one is being transformed (one bulk prefetch, no registers)
    if (t == 0 && img + gridDim.x < p.n_images) {
      const double* nx = p.src + (size_t)(img + gridDim.x) * src_sz;
      const unsigned bytes = (unsigned)((src_sz * sizeof(double)) & ~(size_t)15);
      if ((reinterpret_cast<size_t>(nx) & 15) == 0 && bytes >= 16)   // bulk operations need 16-byte alignment
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(nx), "r"(bytes) : "memory");
    }

    // ---- pass 1: vertical forward radix-12, global -> Z (image columns `line` and `line + 54`)
#pragma unroll 1
    for (int rnd = 0; rnd < 2; ++rnd) {
      const int line = rnd * A_LINES + la;
      if (act && line < HALF) {
        const bool c0 = line < p.ws, c1 = line + HALF < p.ws;
        const double* xr = x + j * p.ws + line;
        const int step = 9 * p.ws;
        cplx v[12];
#pragma unroll
        for (int q = 0; q < 12; ++q) {
          const bool in = j + 9 * q < p.hs;
          v[q].x = (in && c0) ? __ldg(xr + q * step) : 0.0;
          v[q].y = (in && c1) ? __ldg(xr + q * step + HALF) : 0.0;
        }
        bfly12(v);
        cplx* z = Z + line * LD + j;
        z[0] = v[0];
#pragma unroll
        for (int q = 1; q < 12; ++q) z[9 * q] = cmul(v[q], tw[q]);
      }
    }
    __syncthreads();

    // ---- pass 2: vertical forward radix-9 (task = block pb of line `line`)
#pragma unroll 1
    for (int task = t; task < 12 * BPAD; task += THREADS) {
      const int pb = task / BPAD, line = task - pb * BPAD;
      if (line >= HALF) continue;
      cplx* z = Z + line * LD + 9 * pb;
      cplx v[9];
#pragma unroll
      for (int s = 0; s < 9; ++s) v[s] = z[s];
      bfly9(v);
#pragma unroll
      for (int s = 0; s < 9; ++s) z[s] = v[s];
    }
    __syncthreads();

    // ---- pass 3: horizontal, Hermitian unpack + forward radix-12
#pragma unroll 1
    for (int rnd = 0; rnd < 2; ++rnd) {
      const int kk = rnd * A_LINES + la;
      if (act && kk < NK) {
        const int rep = rep_of(kk), par = par_of(kk);
        const bool self = (kk == 0) || (kk == 54);
        cplx* zr = Z + j * LD + rep;
        cplx* zp = Z + j * LD + par;
        cplx v[12];
#pragma unroll
        for (int q = 0; q < 6; ++q) {
          const cplx z1 = zr[9 * q * LD];
          const cplx z2 = self ? z1 : zp[9 * q * LD];
          v[q] = make_double2(z1.x + z2.x, z1.y - z2.y);          // 2 * FFT(column c)      at this frequency
          v[q + 6] = make_double2(z1.y + z2.y, z2.x - z1.x);      // 2 * FFT(column c + 54)
        }
        bfly12(v);
        zr[0] = v[0];
#pragma unroll
        for (int q = 1; q < 6; ++q) zr[9 * q * LD] = cmul(v[q], tw[q]);
#pragma unroll
        for (int q = 6; q < 12; ++q) zp[9 * (q - 6) * LD] = cmul(v[q], tw[q]);
      }
    }
    __syncthreads();

    // ---- pass 4: horizontal radix-9 forward, spectrum product, radix-9 backward
#pragma unroll 1
    for (int task = t; task < 12 * BPAD; task += THREADS) {
      const int pb = task / BPAD, kk = task - pb * BPAD;
      if (kk >= NK) continue;
      const int col = pb < 6 ? rep_of(kk) : par_of(kk);
      cplx* z = Z + (9 * (pb < 6 ? pb : pb - 6)) * LD + col;
      const cplx* __restrict__ ks = p.kspec + (9 * pb) * KLD + kk;
      cplx v[9];
#pragma unroll
      for (int s = 0; s < 9; ++s) v[s] = z[s * LD];
      bfly9(v);
#pragma unroll
      for (int s = 0; s < 9; ++s) v[s] = cswap(cmul(v[s], __ldg(ks + s * KLD)));   // product, swapped for the inverse
      bfly9(v);
#pragma unroll
      for (int s = 0; s < 9; ++s) z[s * LD] = cswap(v[s]);
    }
    __syncthreads();

    // ---- pass 5: horizontal backward radix-12, repack two real columns into one complex line
#pragma unroll 1
    for (int rnd = 0; rnd < 2; ++rnd) {
      const int kk = rnd * A_LINES + la;
      if (act && kk < NK) {
        const int rep = rep_of(kk), par = par_of(kk);
        const bool self = (kk == 0) || (kk == 54);
        cplx* zr = Z + j * LD + rep;
        cplx* zp = Z + j * LD + par;
        cplx v[12];
        v[0] = cswap(zr[0]);
#pragma unroll
        for (int q = 1; q < 6; ++q) v[q] = cswap(cmulc(zr[9 * q * LD], tw[q]));
#pragma unroll
        for (int q = 6; q < 12; ++q) v[q] = cswap(cmulc(zp[9 * (q - 6) * LD], tw[q]));
        bfly12(v);
        swap_all<12>(v);
#pragma unroll
        for (int q = 0; q < 6; ++q) {
          const cplx y1 = v[q], y2 = v[q + 6];
          zr[9 * q * LD] = make_double2(y1.x - y2.y, y1.y + y2.x);              // Y1 + i*Y2 at +freq
          if (!self) zp[9 * q * LD] = make_double2(y1.x + y2.y, y2.x - y1.y);   // conj(Y1) + i*conj(Y2) at -freq
        }
      }
    }
    __syncthreads();

    // ---- pass 6: vertical backward radix-9
#pragma unroll 1
    for (int task = t; task < 12 * BPAD; task += THREADS) {
      const int pb = task / BPAD, line = task - pb * BPAD;
      if (line >= HALF) continue;
      cplx* z = Z + line * LD + 9 * pb;
      cplx v[9];
#pragma unroll
      for (int s = 0; s < 9; ++s) v[s] = cswap(z[s]);
      bfly9(v);
#pragma unroll
      for (int s = 0; s < 9; ++s) z[s] = cswap(v[s]);
    }
    __syncthreads();

    // ---- pass 7: vertical backward radix-12, crop, Z -> global
#pragma unroll 1
    for (int rnd = 0; rnd < 2; ++rnd) {
      const int line = rnd * A_LINES + la;
      if (act && line < HALF) {
        const cplx* z = Z + line * LD + j;
        cplx v[12];
        v[0] = cswap(z[0]);
#pragma unroll
        for (int q = 1; q < 12; ++q) v[q] = cswap(cmulc(z[9 * q], tw[q]));
        bfly12(v);   // results: re/im swapped => v[q].y is column `line`, v[q].x is column `line + 54`
        const int j0 = line - p.off_w, j1 = line + HALF - p.off_w;
        const bool o0 = j0 >= 0 && j0 < p.wd, o1 = j1 >= 0 && j1 < p.wd;
        const int i0 = j - p.off_h;
        double* yr = y + i0 * p.wd + j0;
        const int step = 9 * p.wd;
#pragma unroll
        for (int q = 0; q < 12; ++q) {
          const unsigned i = (unsigned)(i0 + 9 * q);
          const bool in = i < (unsigned)p.hd;
          if (in && o0) yr[q * step] = v[q].y;
          if (in && o1) yr[q * step + HALF] = v[q].x;
        }
      }
    }
    // no barrier: the next image's pass 1 writes exactly the slots this thread just read
  }
}

// Kernel spectrum in the layout pass 4 wants: K[pos][kk] = 0.5/(108*108) * DFT2(kernel)[u(kk)][v(pos)]
// (u = vertical frequency of half-spectrum column kk, v = horizontal frequency at scrambled
// position pos), computed separably straight from the taps (the kernel is tiny; one CTA).
__global__ void __launch_bounds__(512, 1) kspec_kernel(const double* __restrict__ ker, int hk, int wk,
                                                       const cplx* __restrict__ tw, cplx* __restrict__ out) {
  cplx* T1 = reinterpret_cast<cplx*>(b2c_smem_t108);  // [wke][NK]
  const int hke = min(hk, P), wke = min(wk, P);
  for (int idx = threadIdx.x; idx < wke * NK; idx += blockDim.x) {
    const int b = idx / NK, kk = idx - b * NK;
    const int u = freq_of_pos(rep_of(kk));
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
    const int v = freq_of_pos(pos);
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

}  // namespace t108

size_t tuned108_smem_bytes() { return sizeof(cplx) * t108::HALF * t108::LD; }
size_t tuned108_kspec_bytes() { return sizeof(cplx) * t108::KLD * t108::P; }

cudaError_t launch_tuned108_kspec(const double* d_ker, int hk, int wk, const cplx* tw, cplx* out, cudaStream_t st) {
  const size_t smem = sizeof(cplx) * (size_t)(wk < t108::P ? wk : t108::P) * t108::NK;
  cudaFuncSetAttribute(t108::kspec_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  t108::kspec_kernel<<<1, 512, smem, st>>>(d_ker, hk, wk, tw, out);
  return cudaGetLastError();
}

int tuned108_occupancy() {
  int n = 0;
  cudaFuncSetAttribute(t108::conv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tuned108_smem_bytes());
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, t108::conv_kernel, t108::THREADS, tuned108_smem_bytes());
  return n;
}

cudaError_t launch_tuned108_conv(int hs, int ws, int hd, int wd, int off_h, int off_w, const double* d_src,
                                 double* d_dst, long n, const cplx* kspec, const cplx* tw, int grid,
                                 cudaStream_t st) {
  t108::Params p;
  p.hs = hs; p.ws = ws; p.hd = hd; p.wd = wd; p.off_h = off_h; p.off_w = off_w;
  p.src = d_src; p.dst = d_dst; p.n_images = n; p.kspec = kspec; p.tw = tw;
  cudaFuncSetAttribute(t108::conv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tuned108_smem_bytes());
  t108::conv_kernel<<<grid, t108::THREADS, tuned108_smem_bytes(), st>>>(p);
  return cudaGetLastError();
}

}  // namespace b2c
