// Generic fused FFT-convolution kernels: any mode, any 2/3/5/7-smooth period pair, one CTA per
// work buffer.  Replaces fftw_circular_convolution + the crop of convolve
// (reference Convolution/src/convolution_fftw.h:169-284) with ONE kernel per batch:
//   zero-pad on load -> forward rows -> forward columns -> spectrum product (1/N folded in)
//   -> backward columns -> backward rows (only the rows that are kept) -> crop / fold on store.
// Nothing but the source images and the results touch HBM (plus the tiny cached kernel spectrum).
//
// Shared kernel   : two images ride one complex transform, z = a + i*b.  The kernel is real, the
//                   transform linear, so Re/Im of the result are the two convolutions.
// Per-image kernel: z = src + i*kernel in one transform; the two spectra are separated by
//                   Hermitian symmetry inside the product (the reference's GSL back end uses the
//                   same identity, convolution_gsl.h:174-206).
#include "launch_cache.h"
#include "conv_params.h"

namespace b2c {

extern __shared__ __align__(16) unsigned char b2c_smem[];

template <bool SMEM>
__device__ __forceinline__ cplx* work_buffer(const ConvParams& p) {
  if (SMEM) return reinterpret_cast<cplx*>(b2c_smem);
  return p.work + (size_t)blockIdx.x * (size_t)p.ah.P * (size_t)p.ld;
}

// Wrap-accumulated kernel tap at wrapped position (r,c): sum of all taps congruent to it
// (convolution_fftw.h:185-187 builds the same periodic signal with `+=`).
__device__ __forceinline__ double wrapped_tap(const ConvParams& p, const double* ker, int r, int c) {
  double acc = 0.0;
  for (int i = r; i < p.hk; i += p.kwrap_h)
    for (int j = c; j < p.wk; j += p.kwrap_w) acc += ker[(size_t)i * p.wk + j];
  return acc;
}

// ---------------------------------------------------------------------------------------------
// Kernel spectrum: K^ = FFT2(wrap-padded kernel) / (P_h*P_w), scrambled order, to global scratch.
// One CTA.  Replaces the per-call fftw_execute(p_forw_kernel) (convolution_fftw.h:191): with a
// shared kernel it is computed once per batch instead of once per image.
template <bool SMEM>
__global__ void __launch_bounds__(512, 1) kspec_kernel(const __grid_constant__ ConvParams p) {
  cplx* W = work_buffer<SMEM>(p);
  const int Ph = p.ah.P, Pw = p.aw.P, ld = p.ld;
  const float inv_ld = 1.0f / (float)ld, inv_pw = 1.0f / (float)Pw, inv_wd = 1.0f / (float)(p.wd > 0 ? p.wd : 1);
  const int hke = min(p.hk, p.kwrap_h), wke = min(p.wk, p.kwrap_w);
  for (int idx = threadIdx.x; idx < Ph * ld; idx += blockDim.x) {
    const int r = fast_div(idx, ld, inv_ld), c = idx - r * ld;
    double v = 0.0;
    if (r < hke && c < wke) v = wrapped_tap(p, p.ker, r, c);
    W[idx] = make_double2(v, 0.0);
  }
  __syncthreads();
  fft_forward(p.aw, W, 1, ld, hke, false);
  fft_forward(p.ah, W, ld, 1, Pw, true);
  for (int idx = threadIdx.x; idx < Ph * Pw; idx += blockDim.x) {
    const int r = fast_div(idx, Pw, inv_pw), c = idx - r * Pw;
    cplx v = W[r * ld + c];
    p.kspec[idx] = make_double2(v.x * p.scale, v.y * p.scale);
  }
}

// Sum of the (up to four) entries of the linear result that fold onto output (i,j).
__device__ __forceinline__ cplx folded_value(const ConvParams& p, const cplx* W, int i, int j) {
  const int r0 = p.off_h + i, c0 = p.off_w + j;
  cplx v = W[r0 * p.ld + c0];
  if (p.fold_h | p.fold_w) {
    const int r1 = r0 + p.fold_h, c1 = c0 + p.fold_w;
    const bool hr = p.fold_h > 0 && r1 < p.lin_h;
    const bool hc = p.fold_w > 0 && c1 < p.lin_w;
    if (hr) v = cadd(v, W[r1 * p.ld + c0]);
    if (hc) v = cadd(v, W[r0 * p.ld + c1]);
    if (hr && hc) v = cadd(v, W[r1 * p.ld + c1]);
  }
  return v;
}

// ---------------------------------------------------------------------------------------------
// Shared-kernel batch: persistent CTAs, each iteration convolves one PAIR of images.
template <bool SMEM>
__global__ void __launch_bounds__(512, 1) conv_shared_kernel(const __grid_constant__ ConvParams p) {
  cplx* W = work_buffer<SMEM>(p);
  const int Ph = p.ah.P, Pw = p.aw.P, ld = p.ld;
  const float inv_ld = 1.0f / (float)ld, inv_pw = 1.0f / (float)Pw, inv_wd = 1.0f / (float)(p.wd > 0 ? p.wd : 1);
  const long npairs = (p.n_images + 1) / 2;
  const size_t src_sz = (size_t)p.hs * p.ws, dst_sz = (size_t)p.hd * p.wd;
  const int row0 = p.fold_h ? 0 : p.off_h;
  const int nrows = p.fold_h ? p.lin_h : p.hd;

  for (long pair = blockIdx.x; pair < npairs; pair += gridDim.x) {
    const double* a = p.src + (size_t)(2 * pair) * src_sz;
    const bool has_b = 2 * pair + 1 < p.n_images;
    const double* b = a + src_sz;
    // two images ride one transform: image b is scaled by the power of two that equalises the pair's magnitudes
    // (fft_device.cuh: cta_balance) so that the larger image's rounding noise does not land on the smaller one
    const double beta = has_b ? cta_balance(a, (int)src_sz, b, (int)src_sz, reinterpret_cast<double*>(W)) : 1.0;
    const double unbeta = 1.0 / beta;
    // zero-pad on load (convolution_fftw.h:174-183; the source never wraps because P >= n_src)
    // (four elements per thread per trip: eight independent global loads in flight before the stores)
    for (int base = threadIdx.x; base < Ph * ld; base += 4 * blockDim.x) {
      cplx v[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int idx = base + k * blockDim.x;
        const int r = fast_div(idx, ld, inv_ld), c = idx - r * ld;
        v[k] = make_double2(0.0, 0.0);
        if (idx < Ph * ld && r < p.hs && c < p.ws) {
          v[k].x = __ldg(&a[r * p.ws + c]);
          if (has_b) v[k].y = beta * __ldg(&b[r * p.ws + c]);
        }
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int idx = base + k * blockDim.x;
        if (idx < Ph * ld) W[idx] = v[k];
      }
    }
    __syncthreads();
    fft_forward(p.aw, W, 1, ld, p.hs, false);   // rows that hold data
    fft_forward(p.ah, W, ld, 1, Pw, true);      // every column
    for (int idx = threadIdx.x; idx < Ph * Pw; idx += blockDim.x) {   // convolution_fftw.h:195-204
      const int r = fast_div(idx, Pw, inv_pw), c = idx - r * Pw;
      W[r * ld + c] = cmul(W[r * ld + c], __ldg(&p.kspec[idx]));
    }
    __syncthreads();
    fft_backward(p.ah, W, ld, 1, Pw, true);
    fft_backward(p.aw, W + (size_t)row0 * ld, 1, ld, nrows, false);   // only rows that are kept
    // crop / fold on store (convolution_fftw.h:227-272)
    double* da = p.dst + (size_t)(2 * pair) * dst_sz;
    double* db = da + dst_sz;
    for (int idx = threadIdx.x; idx < p.hd * p.wd; idx += blockDim.x) {
      const int i = fast_div(idx, p.wd, inv_wd), j = idx - i * p.wd;
      cplx v = folded_value(p, W, i, j);
      da[idx] = v.x;
      if (has_b) db[idx] = v.y * unbeta;
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------------------------
// Per-image kernels: one image (and its own kernel) per iteration.
template <bool SMEM>
__global__ void __launch_bounds__(512, 1) conv_perimage_kernel(const __grid_constant__ ConvParams p) {
  cplx* W = work_buffer<SMEM>(p);
  const int Ph = p.ah.P, Pw = p.aw.P, ld = p.ld;
  const float inv_ld = 1.0f / (float)ld, inv_pw = 1.0f / (float)Pw, inv_wd = 1.0f / (float)(p.wd > 0 ? p.wd : 1);
  const size_t src_sz = (size_t)p.hs * p.ws, ker_sz = (size_t)p.hk * p.wk, dst_sz = (size_t)p.hd * p.wd;
  const int hke = min(p.hk, p.kwrap_h), wke = min(p.wk, p.kwrap_w);
  const int rows_in = max(p.hs, hke);
  const int row0 = p.fold_h ? 0 : p.off_h;
  const int nrows = p.fold_h ? p.lin_h : p.hd;
  const bool plain_kernel = (p.hk <= p.kwrap_h) && (p.wk <= p.kwrap_w);

  for (long img = blockIdx.x; img < p.n_images; img += gridDim.x) {
    const double* a = p.src + (size_t)img * src_sz;
    const double* k = p.ker + (size_t)img * ker_sz;
    if (threadIdx.x == 0 && img + gridDim.x < p.n_images) {   // next image of this CTA -> L2
      l2_prefetch(p.src + (size_t)(img + gridDim.x) * src_sz, src_sz * sizeof(double));
      l2_prefetch(p.ker + (size_t)(img + gridDim.x) * ker_sz, ker_sz * sizeof(double));
    }
    // kernel part scaled to the source's magnitude (fft_device.cuh: cta_balance), undone in q
    const double al = cta_balance(a, (int)src_sz, k, (int)ker_sz, reinterpret_cast<double*>(W));
    for (int base = threadIdx.x; base < Ph * ld; base += 4 * blockDim.x) {
      cplx v[4];
#pragma unroll
      for (int kq = 0; kq < 4; ++kq) {
        const int idx = base + kq * blockDim.x;
        const int r = fast_div(idx, ld, inv_ld), c = idx - r * ld;
        v[kq] = make_double2(0.0, 0.0);
        if (idx < Ph * ld) {
          if (r < p.hs && c < p.ws) v[kq].x = __ldg(&a[r * p.ws + c]);
          if (r < hke && c < wke) v[kq].y = al * (plain_kernel ? __ldg(&k[r * p.wk + c]) : wrapped_tap(p, k, r, c));
        }
      }
#pragma unroll
      for (int kq = 0; kq < 4; ++kq) {
        const int idx = base + kq * blockDim.x;
        if (idx < Ph * ld) W[idx] = v[kq];
      }
    }
    __syncthreads();
    fft_forward(p.aw, W, 1, ld, rows_in, false);
    fft_forward(p.ah, W, ld, 1, Pw, true);
    // Z = S + iK.  S(w) = (Z(w)+conj Z(-w))/2, K(w) = (Z(w)-conj Z(-w))/(2i)
    // => S*K = (Z(w)^2 - conj(Z(-w))^2) / (4i).  The (-w) partner sits at (neg_h[r], neg_w[c]).
    const double q = 0.25 * p.scale / al;
    for (int idx = threadIdx.x; idx < Ph * Pw; idx += blockDim.x) {
      const int r = fast_div(idx, Pw, inv_pw), c = idx - r * Pw;
      const int pr = __ldg(&p.ah.neg[r]), pc = __ldg(&p.aw.neg[c]);
      const int pidx = pr * Pw + pc;
      if (pidx < idx) continue;  // the partner's thread writes both
      cplx z1 = W[r * ld + c], z2 = W[pr * ld + pc];
      cplx d = make_double2(z1.x * z1.x - z1.y * z1.y - (z2.x * z2.x - z2.y * z2.y),
                            2.0 * (z1.x * z1.y + z2.x * z2.y));      // z1^2 - conj(z2)^2
      cplx y = make_double2(d.y * q, -d.x * q);                       // * (-i/4) * scale
      W[r * ld + c] = y;
      if (pidx != idx) W[pr * ld + pc] = make_double2(y.x, -y.y);
    }
    __syncthreads();
    fft_backward(p.ah, W, ld, 1, Pw, true);
    fft_backward(p.aw, W + (size_t)row0 * ld, 1, ld, nrows, false);
    double* d = p.dst + (size_t)img * dst_sz;
    for (int idx = threadIdx.x; idx < p.hd * p.wd; idx += blockDim.x) {
      const int i = fast_div(idx, p.wd, inv_wd), j = idx - i * p.wd;
      d[idx] = folded_value(p, W, i, j).x;
    }
    __syncthreads();
  }
}

// ---- host-visible launchers -------------------------------------------------------------------
cudaError_t launch_kspec(const ConvParams& p, int threads, size_t smem, cudaStream_t st) {
  if (p.work_in_smem) {
    ensure_dyn_smem(reinterpret_cast<const void*>(kspec_kernel<true>), smem);
    kspec_kernel<true><<<1, threads, smem, st>>>(p);
  } else {
    kspec_kernel<false><<<1, threads, 0, st>>>(p);
  }
  return cudaGetLastError();
}

cudaError_t launch_conv_shared(const ConvParams& p, int grid, int threads, size_t smem, cudaStream_t st) {
  if (p.work_in_smem) {
    ensure_dyn_smem(reinterpret_cast<const void*>(conv_shared_kernel<true>), smem);
    conv_shared_kernel<true><<<grid, threads, smem, st>>>(p);
  } else {
    conv_shared_kernel<false><<<grid, threads, 0, st>>>(p);
  }
  return cudaGetLastError();
}

cudaError_t launch_conv_perimage(const ConvParams& p, int grid, int threads, size_t smem, cudaStream_t st) {
  if (p.work_in_smem) {
    ensure_dyn_smem(reinterpret_cast<const void*>(conv_perimage_kernel<true>), smem);
    conv_perimage_kernel<true><<<grid, threads, smem, st>>>(p);
  } else {
    conv_perimage_kernel<false><<<grid, threads, 0, st>>>(p);
  }
  return cudaGetLastError();
}

int occupancy_conv_shared(int threads, size_t smem, bool in_smem) {
  int n = 0;
  if (in_smem) {
    ensure_dyn_smem(reinterpret_cast<const void*>(conv_shared_kernel<true>), smem);
    n = cached_occupancy(reinterpret_cast<const void*>(conv_shared_kernel<true>), threads, smem);
  } else {
    n = cached_occupancy(reinterpret_cast<const void*>(conv_shared_kernel<false>), threads, 0);
  }
  return n;
}

int occupancy_conv_perimage(int threads, size_t smem, bool in_smem) {
  int n = 0;
  if (in_smem) {
    ensure_dyn_smem(reinterpret_cast<const void*>(conv_perimage_kernel<true>), smem);
    n = cached_occupancy(reinterpret_cast<const void*>(conv_perimage_kernel<true>), threads, smem);
  } else {
    n = cached_occupancy(reinterpret_cast<const void*>(conv_perimage_kernel<false>), threads, 0);
  }
  return n;
}

}  // namespace b2c
