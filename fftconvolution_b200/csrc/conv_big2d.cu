// Fused FFT convolution for transforms whose work buffer does not fit in shared memory
// (e.g. BASELINE config 3: 100x100 (*) 100x100 CIRCULAR_FULL -> period 200 x 200 = 640 KB complex).
// Replaces fftw_circular_convolution + crop (reference convolution_fftw.h:169-284) for such plans.
//
// One persistent CTA per SM owns an image and walks three phases (148 slabs of ~0.65 MB stay inside
// the 126 MB L2; two CTAs per SM measured no faster and spill the slabs to HBM); inside a phase every transform
// runs entirely in shared memory, between phases only HALF spectra (real input) travel through a
// per-CTA global slab that stays L2-resident:
//   R  : chunks of rows.  Source row r and kernel row r ride one complex transform
//        (z = s_r + i*k_r); the two half spectra are separated by Hermitian symmetry and written to
//        the slab                                                        [1 slab write]
//   C  : strips of half-spectrum columns.  Column transforms of both spectra, pointwise product
//        (1/N folded in), backward column transform                      [1 slab read, 1 slab write]
//   R' : chunks of row pairs.  Rows r and r+1 of the result ride one backward transform
//        (W = Y_r + i*Y_{r+1}, Hermitian-extended), crop on store        [1 slab read]
// The any-size kernel of conv_generic.cu would instead push every radix pass of the full complex
// transform through the slab (~24 slab passes); this schedule needs 4 (the result Y replaces the S|K
// columns of its strip in place, so a slab is one half spectrum of the padded image pair).
// Two forms share the phase code: the persistent one above for batches that fill the machine, and a
// split one (one launch per phase, one CTA per chunk / strip, one slab per image) so that a few large
// images -- down to a single 2048 x 2048 one -- still use every SM.
#include "conv_params.h"

namespace b2c {

extern __shared__ __align__(16) unsigned char b2c_smem_big[];

__device__ __forceinline__ double big_wrapped_tap(const BigParams& p, const double* ker, int r, int c) {
  double acc = 0.0;   // taps congruent modulo the period add up (convolution_fftw.h:185-187)
  for (int i = r; i < p.hk; i += p.kwrap_h)
    for (int j = c; j < p.wk; j += p.kwrap_w) acc += ker[(size_t)i * p.wk + j];
  return acc;
}

// Everything a phase needs besides its chunk / strip index.
struct BigCtx {
  cplx* W;
  int Ph, Pw, hke, wke, rows_in, HW;
  bool plain_kernel;
  float inv_ldw, inv_hw, inv_ldc, inv_wd, inv_pw;
  cplx *slabS, *slabK, *slabY;
  double alpha;   // power-of-two scale of the kernel rows in the packed row transforms (balance_kernel)
};

__device__ __forceinline__ BigCtx big_ctx(const BigParams& p, size_t slab_index, long img) {
  BigCtx c;
  c.alpha = img < p.n_images ? __ldg(p.alpha + img) : 1.0;
  c.W = reinterpret_cast<cplx*>(b2c_smem_big);
  c.Ph = p.ah.P; c.Pw = p.aw.P;
  c.hke = min(p.hk, p.kwrap_h); c.wke = min(p.wk, p.kwrap_w);
  c.rows_in = max(p.hs, c.hke);
  c.HW = p.half_w;
  c.plain_kernel = (p.hk <= p.kwrap_h) && (p.wk <= p.kwrap_w);
  c.inv_ldw = 1.0f / (float)p.ldw; c.inv_hw = 1.0f / (float)c.HW; c.inv_ldc = 1.0f / (float)p.ldc;
  c.inv_wd = 1.0f / (float)p.wd; c.inv_pw = 1.0f / (float)c.Pw;
  c.slabS = p.slab + slab_index * (size_t)max(p.hs + c.hke, p.rows_y) * c.HW;
  c.slabK = c.slabS + (size_t)p.hs * c.HW;
  c.slabY = c.slabS;   // in place: a strip's Y columns replace the S and K columns it consumed
  return c;
}

// ---------------- phase R, one chunk of rows: forward row transforms, Hermitian split, -> slab
__device__ __forceinline__ void big_phase_r(const BigParams& p, const BigCtx& x, const double* a, const double* k, int r0) {
  cplx* W = x.W;
  const int nr = min(p.rows_chunk, x.rows_in - r0);
  const int nelem = nr * p.ldw;
  for (int base = threadIdx.x; base < nelem; base += 4 * blockDim.x) {
    cplx v[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int idx = base + q * blockDim.x;
      const int rr = fast_div(idx, p.ldw, x.inv_ldw), c = idx - rr * p.ldw;
      const int r = r0 + rr;
      v[q] = make_double2(0.0, 0.0);
      if (idx < nelem) {
        if (r < p.hs && c < p.ws) v[q].x = __ldg(&a[(size_t)r * p.ws + c]);
        if (r < x.hke && c < x.wke) v[q].y = x.alpha * (x.plain_kernel ? __ldg(&k[(size_t)r * p.wk + c]) : big_wrapped_tap(p, k, r, c));
      }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int idx = base + q * blockDim.x;
      if (idx < nelem) W[idx] = v[q];
    }
  }
  __syncthreads();
  fft_forward(p.aw, W, 1, p.ldw, nr, false);
  // Z = S + iK  =>  S(v) = (Z(v) + conj Z(-v)) / 2,  K(v) = -i (Z(v) - conj Z(-v)) / 2
  const int HW = x.HW;
  for (int idx = threadIdx.x; idx < nr * HW; idx += blockDim.x) {
    const int rr = fast_div(idx, HW, x.inv_hw), v = idx - rr * HW;
    const int r = r0 + rr;
    const cplx z1 = W[rr * p.ldw + __ldg(&p.aw.posf[v])];
    const cplx z2 = W[rr * p.ldw + __ldg(&p.aw.posf[v ? x.Pw - v : 0])];
    if (r < p.hs) x.slabS[(size_t)r * HW + v] = make_double2(0.5 * (z1.x + z2.x), 0.5 * (z1.y - z2.y));
    if (r < x.hke) x.slabK[(size_t)r * HW + v] = make_double2(0.5 * (z1.y + z2.y), 0.5 * (z2.x - z1.x));
  }
  __syncthreads();
}

// ---------------- phase C, one strip of half-spectrum columns: column transforms of both spectra,
// product, backward columns
__device__ __forceinline__ void big_phase_c(const BigParams& p, const BigCtx& x, int v0) {
  cplx* W = x.W;
  const int HW = x.HW, Ph = x.Ph;
  const int nc = min(p.cols_strip, HW - v0);
  const int nelem = Ph * p.ldc;
  for (int base = threadIdx.x; base < nelem; base += 4 * blockDim.x) {
    cplx v[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int idx = base + q * blockDim.x;
      const int r = fast_div(idx, p.ldc, x.inv_ldc), c = idx - r * p.ldc;
      v[q] = make_double2(0.0, 0.0);
      if (idx < nelem) {
        if (c < nc) { if (r < p.hs) v[q] = x.slabS[(size_t)r * HW + v0 + c]; }
        else if (c < 2 * nc) { if (r < x.hke) v[q] = x.slabK[(size_t)r * HW + v0 + c - nc]; }
      }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int idx = base + q * blockDim.x;
      if (idx < nelem) W[idx] = v[q];
    }
  }
  __syncthreads();
  fft_forward(p.ah, W, p.ldc, 1, 2 * nc, true);
  const double sc = p.scale / x.alpha;
  for (int idx = threadIdx.x; idx < Ph * nc; idx += blockDim.x) {   // convolution_fftw.h:195-204, 210-211
    const int r = idx / nc, c = idx - r * nc;
    const cplx s = W[r * p.ldc + c], kk = W[r * p.ldc + nc + c];
    const cplx y = cmul(s, kk);
    W[r * p.ldc + c] = make_double2(y.x * sc, y.y * sc);
  }
  __syncthreads();
  fft_backward(p.ah, W, p.ldc, 1, nc, true);
  for (int idx = threadIdx.x; idx < p.rows_y * nc; idx += blockDim.x) {
    const int r = idx / nc, c = idx - r * nc;
    x.slabY[(size_t)r * HW + v0 + c] = W[r * p.ldc + c];
  }
  __syncthreads();
}

// ---------------- phase R', one chunk of row pairs: backward row transforms, two result rows per
// transform, crop, -> dst
__device__ __forceinline__ void big_phase_rp(const BigParams& p, const BigCtx& x, double* d, int q0) {
  cplx* W = x.W;
  const int HW = x.HW, Pw = x.Pw;
  const int pairs = (p.hd + 1) / 2;
  const int nq = min(p.rows_chunk, pairs - q0);
  const int nelem = nq * Pw;
  for (int base = threadIdx.x; base < nelem; base += 4 * blockDim.x) {   // 8 slab loads in flight per thread
    cplx y0[4], y1[4];
    int dstpos[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int idx = base + u * blockDim.x;
      dstpos[u] = -1;
      if (idx < nelem) {
        const int qq = fast_div(idx, Pw, x.inv_pw), v = idx - qq * Pw;
        const int i0 = 2 * (q0 + qq), i1 = i0 + 1;         // output rows
        const int vv = v < HW ? v : Pw - v;
        y0[u] = x.slabY[(size_t)(p.off_h + i0) * HW + vv];
        y1[u] = i1 < p.hd ? x.slabY[(size_t)(p.off_h + i1) * HW + vv] : make_double2(0.0, 0.0);
        if (v >= HW) { y0[u].y = -y0[u].y; y1[u].y = -y1[u].y; }        // Hermitian mirror
        dstpos[u] = qq * p.ldw + __ldg(&p.aw.posf[v]);
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (dstpos[u] >= 0) W[dstpos[u]] = make_double2(y0[u].x - y1[u].y, y0[u].y + y1[u].x);   // Y0 + i*Y1
  }
  __syncthreads();
  fft_backward(p.aw, W, 1, p.ldw, nq, false);
  for (int idx = threadIdx.x; idx < 2 * nq * p.wd; idx += blockDim.x) {
    const int ri = fast_div(idx, p.wd, x.inv_wd), j = idx - ri * p.wd;
    const int i = 2 * q0 + ri;
    if (i < p.hd) {
      const cplx w = W[(ri >> 1) * p.ldw + p.off_w + j];
      d[(size_t)i * p.wd + j] = (ri & 1) ? w.y : w.x;
    }
  }
  __syncthreads();
}

// Persistent form: one CTA per SM owns an image and walks the three phases through its own slab.
__global__ void __launch_bounds__(512, 1) conv_big2d_kernel(const __grid_constant__ BigParams p) {
  BigCtx x = big_ctx(p, blockIdx.x, blockIdx.x);
  const size_t src_sz = (size_t)p.hs * p.ws, dst_sz = (size_t)p.hd * p.wd;
  const int pairs = (p.hd + 1) / 2;
  for (long img = blockIdx.x; img < p.n_images; img += gridDim.x) {
    x.alpha = __ldg(p.alpha + img);
    const double* a = p.src + (size_t)img * src_sz;
    const double* k = p.ker + (size_t)img * p.ker_stride;
    for (int r0 = 0; r0 < x.rows_in; r0 += p.rows_chunk) big_phase_r(p, x, a, k, r0);
    for (int v0 = 0; v0 < x.HW; v0 += p.cols_strip) big_phase_c(p, x, v0);
    double* d = p.dst + (size_t)img * dst_sz;
    for (int q0 = 0; q0 < pairs; q0 += p.rows_chunk) big_phase_rp(p, x, d, q0);
  }
}

// Split form for batches smaller than the machine (large images, few of them): one launch per phase,
// one CTA per (image, chunk or strip), one slab per IMAGE -- every SM works on the same image.
template <int PHASE>
__global__ void __launch_bounds__(512, 1) conv_big2d_phase_kernel(const __grid_constant__ BigParams p, int units) {
  const long img = blockIdx.x / units;
  const int unit = blockIdx.x - (int)(img * units);
  const BigCtx x = big_ctx(p, (size_t)img, img);
  if (PHASE == 0) {
    big_phase_r(p, x, p.src + (size_t)img * p.hs * p.ws, p.ker + (size_t)img * p.ker_stride, unit * p.rows_chunk);
  } else if (PHASE == 1) {
    big_phase_c(p, x, unit * p.cols_strip);
  } else {
    big_phase_rp(p, x, p.dst + (size_t)img * p.hd * p.wd, unit * p.rows_chunk);
  }
}

cudaError_t launch_conv_big2d(const BigParams& p, int grid, size_t smem, cudaStream_t st) {
  cudaFuncSetAttribute(conv_big2d_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  conv_big2d_kernel<<<grid, 512, smem, st>>>(p);
  return cudaGetLastError();
}

// p.slab holds one slab per image of this launch group (p.n_images of them).
cudaError_t launch_conv_big2d_split(const BigParams& p, size_t smem, cudaStream_t st, int* launches) {
  const int hke = p.hk < p.kwrap_h ? p.hk : p.kwrap_h;
  const int rows_in = p.hs > hke ? p.hs : hke, pairs = (p.hd + 1) / 2;
  const int ur = (rows_in + p.rows_chunk - 1) / p.rows_chunk;
  const int uc = (p.half_w + p.cols_strip - 1) / p.cols_strip;
  const int up = (pairs + p.rows_chunk - 1) / p.rows_chunk;
  cudaFuncSetAttribute(conv_big2d_phase_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaFuncSetAttribute(conv_big2d_phase_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaFuncSetAttribute(conv_big2d_phase_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  conv_big2d_phase_kernel<0><<<(unsigned)(p.n_images * ur), 512, smem, st>>>(p, ur);
  conv_big2d_phase_kernel<1><<<(unsigned)(p.n_images * uc), 512, smem, st>>>(p, uc);
  conv_big2d_phase_kernel<2><<<(unsigned)(p.n_images * up), 512, smem, st>>>(p, up);
  *launches = 3;
  return cudaGetLastError();
}

}  // namespace b2c
