// Headline kernel: fused convolution at period 108 x 108 with a shared kernel (BASELINE configs[1]:
// 100 x 100 (*) 15 x 15 LINEAR_SAME needs a period >= 107).  Five in-place passes over one shared buffer, prime-factor
// 4 x 27 transforms without inter-stage twiddles -- the passes themselves are in conv_pfa108.cuh.
// One persistent CTA of 256 threads per image slot, two CTAs per SM.
// Replaces fftw_circular_convolution + crop (reference convolution_fftw.h:169-284).
#include "conv_pfa108.cuh"

namespace b2c {

extern __shared__ __align__(16) unsigned char b2c_smem_pfa[];

namespace pfa {

#ifdef B2C_PFA_TRACE
// development variant (tools/trace_pfa108.py): clock stamps of every warp around each pass of a CTA's fourth image
__device__ long long g_trace[512 * 8 * 16];
#define B2C_STAMP(k)                                                                          \
  if (traced && (t & 31) == 0) g_trace[(blockIdx.x * 8 + (t >> 5)) * 16 + (k)] = clock64();
#else
#define B2C_STAMP(k)
#endif

__global__ void __launch_bounds__(THREADS, 2) conv_kernel(const __grid_constant__ Params p) {
  cplx* Z = reinterpret_cast<cplx*>(b2c_smem_pfa);
  const int t = threadIdx.x;
  const size_t src_sz = (size_t)p.hs * p.ws, dst_sz = (size_t)p.hd * p.wd;
  for (long img = blockIdx.x; img < p.n_images; img += gridDim.x) {
#ifdef B2C_PFA_TRACE
    const bool traced = img == blockIdx.x + 3L * gridDim.x && blockIdx.x < 512;
#endif
    B2C_STAMP(0)
    if (t == 0 && img + gridDim.x < p.n_images)   // next image -> L2
      l2_prefetch(p.src + (size_t)(img + gridDim.x) * src_sz, src_sz * sizeof(double));
    pass1(t, Z, p, p.src + (size_t)img * src_sz);
    B2C_STAMP(1)
    __syncthreads();
    B2C_STAMP(2)
    pass2(t, Z);
    B2C_STAMP(3)
    __syncthreads();
    B2C_STAMP(4)
    pass3(t, Z, p.kspec);
    B2C_STAMP(5)
    __syncthreads();
    B2C_STAMP(6)
    pass4(t, Z);
    B2C_STAMP(7)
    __syncthreads();
    B2C_STAMP(8)
    pass5(t, Z, p, p.dst + (size_t)img * dst_sz);
    B2C_STAMP(9)
    // no barrier: the next image's pass 1 writes exactly the slots this thread just read
  }
}

// Kernel spectrum K[k4'][k27'][kk] = 0.5/108^2 * DFT2(kernel)[kv(kk)][kh(k4', k27')], straight from the taps; taps
// congruent modulo the period add up (convolution_fftw.h:185-187).  One CTA, once per batch.
__global__ void __launch_bounds__(512, 1) kspec_kernel(const double* __restrict__ ker, int hk, int wk,
                                                       const cplx* __restrict__ tw, cplx* __restrict__ out) {
  cplx* T1 = reinterpret_cast<cplx*>(b2c_smem_pfa);   // [wke][KK]
  const int hke = min(hk, P), wke = min(wk, P);
  for (int idx = threadIdx.x; idx < wke * KK; idx += blockDim.x) {
    const int b = idx / KK, kk = idx - b * KK;
    const int u = kv_of(kk);
    cplx acc = make_double2(0.0, 0.0);
    for (int a = 0; a < hke; ++a) {
      double tap = 0.0;
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
  for (int idx = threadIdx.x; idx < P * KK; idx += blockDim.x) {
    const int pos = idx / KK, kk = idx - pos * KK;
    const int v = kh_of(pos / 27, pos % 27);
    cplx acc = make_double2(0.0, 0.0);
    for (int b = 0; b < wke; ++b) {
      const cplx a = T1[b * KK + kk];
      const cplx w = tw[(v * b) % P];
      acc.x += a.x * w.x - a.y * w.y;
      acc.y += a.x * w.y + a.y * w.x;
    }
    out[idx] = make_double2(acc.x * scale, acc.y * scale);
  }
}

}  // namespace pfa

#ifdef B2C_PFA_TRACE
extern "C" int b2c_debug_trace_read(long long* out, long n) {
  return (int)cudaMemcpyFromSymbol(out, pfa::g_trace, sizeof(long long) * (size_t)(n < 512 * 8 * 16 ? n : 512 * 8 * 16));
}
#endif

size_t pfa108_kspec_bytes() { return sizeof(cplx) * (size_t)pfa::P * pfa::KK; }

int pfa108_occupancy() {
  int occ = 0;
  cudaFuncSetAttribute(pfa::conv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pfa::SMEM_BYTES);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, pfa::conv_kernel, pfa::THREADS, pfa::SMEM_BYTES);
  return occ;
}

// kernel spectrum (skipped when d_ker is null: the spectrum of the previous call is still valid)
cudaError_t launch_pfa108_kspec(const double* d_ker, int hk, int wk, const cplx* tw, cplx* kspec, cudaStream_t st) {
  const size_t ksmem = sizeof(cplx) * (size_t)(wk < pfa::P ? wk : pfa::P) * pfa::KK;
  cudaFuncSetAttribute(pfa::kspec_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ksmem);
  pfa::kspec_kernel<<<1, 512, ksmem, st>>>(d_ker, hk, wk, tw, kspec);
  return cudaGetLastError();
}

cudaError_t launch_pfa108_conv(int hs, int ws, int hd, int wd, int off_h, int off_w, int hs_in, int ws_in, int sh_h,
                               int sh_w, const double* d_src, double* d_dst, long n, const cplx* kspec, int grid,
                               cudaStream_t st) {
  pfa::Params p;
  p.hs = hs; p.ws = ws; p.hd = hd; p.wd = wd; p.off_h = off_h; p.off_w = off_w;
  p.hs_in = hs_in; p.ws_in = ws_in; p.sh_h = sh_h; p.sh_w = sh_w;
  p.src = d_src; p.dst = d_dst; p.n_images = n; p.kspec = kspec;
  cudaFuncSetAttribute(pfa::conv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pfa::SMEM_BYTES);
  pfa::conv_kernel<<<grid, pfa::THREADS, pfa::SMEM_BYTES, st>>>(p);
  return cudaGetLastError();
}

// kernel spectrum + convolution; *launches += 2 (1 when the spectrum is reused)
cudaError_t launch_pfa108(int hs, int ws, int hk, int wk, int hd, int wd, int off_h, int off_w, int hs_in, int ws_in,
                          int sh_h, int sh_w, const double* d_src, const double* d_ker, double* d_dst, long n,
                          cplx* kspec, const cplx* tw, int sms, cudaStream_t st, int* launches) {
  if (d_ker) {
    cudaError_t e = launch_pfa108_kspec(d_ker, hk, wk, tw, kspec, st);
    if (e != cudaSuccess) return e;
    ++*launches;
  }
  const int occ = pfa108_occupancy();
  if (occ < 1) return cudaErrorLaunchOutOfResources;
  long grid = (long)sms * occ;
  if (grid > n) grid = n;
  ++*launches;
  return launch_pfa108_conv(hs, ws, hd, wd, off_h, off_w, hs_in, ws_in, sh_h, sh_w, d_src, d_dst, n, kspec, (int)grid, st);
}

}  // namespace b2c
