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
// transform through the slab (~24 slab passes); this schedule needs 4.
#include "conv_params.h"

namespace b2c {

extern __shared__ __align__(16) unsigned char b2c_smem_big[];

__device__ __forceinline__ double big_wrapped_tap(const BigParams& p, const double* ker, int r, int c) {
  double acc = 0.0;   // taps congruent modulo the period add up (convolution_fftw.h:185-187)
  for (int i = r; i < p.hk; i += p.kwrap_h)
    for (int j = c; j < p.wk; j += p.kwrap_w) acc += ker[(size_t)i * p.wk + j];
  return acc;
}

__global__ void __launch_bounds__(512, 1) conv_big2d_kernel(const __grid_constant__ BigParams p) {
  cplx* W = reinterpret_cast<cplx*>(b2c_smem_big);
  const int Ph = p.ah.P, Pw = p.aw.P;
  const int hke = min(p.hk, p.kwrap_h), wke = min(p.wk, p.kwrap_w);
  const int rows_in = max(p.hs, hke);
  const int HW = p.half_w;
  const bool plain_kernel = (p.hk <= p.kwrap_h) && (p.wk <= p.kwrap_w);
  const size_t src_sz = (size_t)p.hs * p.ws, dst_sz = (size_t)p.hd * p.wd;
  cplx* slabS = p.slab + (size_t)blockIdx.x * (size_t)max(p.hs + hke, p.rows_y) * HW;
  cplx* slabK = slabS + (size_t)p.hs * HW;
  cplx* slabY = slabS;   // in place: a strip's Y columns replace the S and K columns it consumed
  const float inv_ldw = 1.0f / (float)p.ldw, inv_hw = 1.0f / (float)HW, inv_ldc = 1.0f / (float)p.ldc;
  const float inv_wd = 1.0f / (float)p.wd, inv_pw = 1.0f / (float)Pw;

  for (long img = blockIdx.x; img < p.n_images; img += gridDim.x) {
    const double* a = p.src + (size_t)img * src_sz;
    const double* k = p.ker + (size_t)img * p.ker_stride;

    // ---------------- phase R: forward row transforms, Hermitian split, -> slab
    for (int r0 = 0; r0 < rows_in; r0 += p.rows_chunk) {
      const int nr = min(p.rows_chunk, rows_in - r0);
      const int nelem = nr * p.ldw;
      for (int base = threadIdx.x; base < nelem; base += 4 * blockDim.x) {
        cplx v[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int idx = base + q * blockDim.x;
          const int rr = fast_div(idx, p.ldw, inv_ldw), c = idx - rr * p.ldw;
          const int r = r0 + rr;
          v[q] = make_double2(0.0, 0.0);
          if (idx < nelem) {
            if (r < p.hs && c < p.ws) v[q].x = __ldg(&a[(size_t)r * p.ws + c]);
            if (r < hke && c < wke) v[q].y = plain_kernel ? __ldg(&k[(size_t)r * p.wk + c]) : big_wrapped_tap(p, k, r, c);
          }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int idx = base + q * blockDim.x;
          if (idx < nelem) W[idx] = v[q];
        }
      }
      __syncthreads();
      if (is_static_200(p.aw)) StaticFft<200, 5, 5, 8, 1>::forward(p.aw.tw, W, 1, p.ldw, nr, false);
      else fft_forward(p.aw, W, 1, p.ldw, nr, false);
      // Z = S + iK  =>  S(v) = (Z(v) + conj Z(-v)) / 2,  K(v) = -i (Z(v) - conj Z(-v)) / 2
      for (int idx = threadIdx.x; idx < nr * HW; idx += blockDim.x) {
        const int rr = fast_div(idx, HW, inv_hw), v = idx - rr * HW;
        const int r = r0 + rr;
        const cplx z1 = W[rr * p.ldw + __ldg(&p.aw.posf[v])];
        const cplx z2 = W[rr * p.ldw + __ldg(&p.aw.posf[v ? Pw - v : 0])];
        if (r < p.hs) slabS[(size_t)r * HW + v] = make_double2(0.5 * (z1.x + z2.x), 0.5 * (z1.y - z2.y));
        if (r < hke) slabK[(size_t)r * HW + v] = make_double2(0.5 * (z1.y + z2.y), 0.5 * (z2.x - z1.x));
      }
      __syncthreads();
    }

    // ---------------- phase C: column transforms of both half spectra, product, backward columns
    for (int v0 = 0; v0 < HW; v0 += p.cols_strip) {
      const int nc = min(p.cols_strip, HW - v0);
      const int nelem = Ph * p.ldc;
      for (int base = threadIdx.x; base < nelem; base += 4 * blockDim.x) {
        cplx v[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int idx = base + q * blockDim.x;
          const int r = fast_div(idx, p.ldc, inv_ldc), c = idx - r * p.ldc;
          v[q] = make_double2(0.0, 0.0);
          if (idx < nelem) {
            if (c < nc) { if (r < p.hs) v[q] = slabS[(size_t)r * HW + v0 + c]; }
            else if (c < 2 * nc) { if (r < hke) v[q] = slabK[(size_t)r * HW + v0 + c - nc]; }
          }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int idx = base + q * blockDim.x;
          if (idx < nelem) W[idx] = v[q];
        }
      }
      __syncthreads();
      if (is_static_200(p.ah)) StaticFft<200, 5, 5, 8, 1>::forward(p.ah.tw, W, p.ldc, 1, 2 * nc, true);
      else fft_forward(p.ah, W, p.ldc, 1, 2 * nc, true);
      for (int idx = threadIdx.x; idx < Ph * nc; idx += blockDim.x) {   // convolution_fftw.h:195-204, 210-211
        const int r = idx / nc, c = idx - r * nc;
        const cplx s = W[r * p.ldc + c], kk = W[r * p.ldc + nc + c];
        const cplx y = cmul(s, kk);
        W[r * p.ldc + c] = make_double2(y.x * p.scale, y.y * p.scale);
      }
      __syncthreads();
      if (is_static_200(p.ah)) StaticFft<200, 5, 5, 8, 1>::backward(p.ah.tw, W, p.ldc, 1, nc, true);
      else fft_backward(p.ah, W, p.ldc, 1, nc, true);
      for (int idx = threadIdx.x; idx < p.rows_y * nc; idx += blockDim.x) {
        const int r = idx / nc, c = idx - r * nc;
        slabY[(size_t)r * HW + v0 + c] = W[r * p.ldc + c];
      }
      __syncthreads();
    }

    // ---------------- phase R': backward row transforms, two result rows per transform, crop, -> dst
    double* d = p.dst + (size_t)img * dst_sz;
    const int pairs = (p.hd + 1) / 2;
    for (int q0 = 0; q0 < pairs; q0 += p.rows_chunk) {
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
            const int qq = fast_div(idx, Pw, inv_pw), v = idx - qq * Pw;
            const int i0 = 2 * (q0 + qq), i1 = i0 + 1;         // output rows
            const int vv = v < HW ? v : Pw - v;
            y0[u] = slabY[(size_t)(p.off_h + i0) * HW + vv];
            y1[u] = i1 < p.hd ? slabY[(size_t)(p.off_h + i1) * HW + vv] : make_double2(0.0, 0.0);
            if (v >= HW) { y0[u].y = -y0[u].y; y1[u].y = -y1[u].y; }        // Hermitian mirror
            dstpos[u] = qq * p.ldw + __ldg(&p.aw.posf[v]);
          }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (dstpos[u] >= 0) W[dstpos[u]] = make_double2(y0[u].x - y1[u].y, y0[u].y + y1[u].x);   // Y0 + i*Y1
      }
      __syncthreads();
      if (is_static_200(p.aw)) StaticFft<200, 5, 5, 8, 1>::backward(p.aw.tw, W, 1, p.ldw, nq, false);
      else fft_backward(p.aw, W, 1, p.ldw, nq, false);
      for (int idx = threadIdx.x; idx < 2 * nq * p.wd; idx += blockDim.x) {
        const int ri = fast_div(idx, p.wd, inv_wd), j = idx - ri * p.wd;
        const int i = 2 * q0 + ri;
        if (i < p.hd) {
          const cplx w = W[(ri >> 1) * p.ldw + p.off_w + j];
          d[(size_t)i * p.wd + j] = (ri & 1) ? w.y : w.x;
        }
      }
      __syncthreads();
    }
  }
}

cudaError_t launch_conv_big2d(const BigParams& p, int grid, size_t smem, cudaStream_t st) {
  cudaFuncSetAttribute(conv_big2d_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  conv_big2d_kernel<<<grid, 512, smem, st>>>(p);
  return cudaGetLastError();
}

}  // namespace b2c
