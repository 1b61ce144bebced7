// Direct (nested-loop) convolution kernel for small kernels, where the reference itself finds
// loops faster than the FFT (README.md:31,47; Convolution/src/convolution_std.h:86-199 is the
// CPU form).  One CTA stages a row tile of the source (zero- or wrap-padded halo) plus the
// flipped kernel in shared memory; every thread produces TJ=8 adjacent outputs with a sliding
// register window, so each shared-memory word feeds 8 FP64 FMAs.
#include "conv_params.h"

namespace b2c {

extern __shared__ __align__(16) unsigned char b2c_smem_direct[];

constexpr int TJ = 8;

__device__ __forceinline__ int wrap_index(int x, int m) {
  x %= m;
  return x < 0 ? x + m : x;
}

__global__ void __launch_bounds__(256, 2) direct_kernel(const __grid_constant__ DirectParams p) {
  double* kflip = reinterpret_cast<double*>(b2c_smem_direct);
  const int wk_pad = (p.wk + TJ - 1) / TJ * TJ;
  double* S = kflip + (size_t)p.hk * wk_pad;
  const int sp_w = p.sp_w;
  const size_t src_sz = (size_t)p.hs * p.ws, ker_sz = (size_t)p.hk * p.wk, dst_sz = (size_t)p.hd * p.wd;
  const int groups = (p.wd + TJ - 1) / TJ;
  const long total = p.n_images * p.tiles_per_image;
  bool kernel_loaded = false;

  for (long work = blockIdx.x; work < total; work += gridDim.x) {
    const long img = work / p.tiles_per_image;
    const int tile = (int)(work - img * p.tiles_per_image);
    const int i0 = tile * p.tile_h;
    const int th = min(p.tile_h, p.hd - i0);
    __syncthreads();  // previous iteration's readers are done
    if (!(p.kernel_is_shared && kernel_loaded)) {
      const double* k = p.ker + (p.kernel_is_shared ? 0 : (size_t)img * ker_sz);
      for (int idx = threadIdx.x; idx < p.hk * wk_pad; idx += blockDim.x) {
        const int r = idx / wk_pad, c = idx - r * wk_pad;
        kflip[idx] = c < p.wk ? __ldg(&k[(size_t)(p.hk - 1 - r) * p.wk + (p.wk - 1 - c)]) : 0.0;
      }
      kernel_loaded = true;
    }
    // stage source rows i0+off_h-(hk-1) ... with halo
    const double* a = p.src + (size_t)img * src_sz;
    const int rows = th + p.hk - 1;
    for (int idx = threadIdx.x; idx < rows * sp_w; idx += blockDim.x) {
      const int u = idx / sp_w, v = idx - u * sp_w;
      int sr = i0 + u + p.off_h - (p.hk - 1);
      int sc = v + p.off_w - (p.wk - 1);
      double val = 0.0;
      if (v < p.wd + p.wk - 1) {
        if (p.wrap_h) sr = wrap_index(sr, p.wrap_h);
        if (p.wrap_w) sc = wrap_index(sc, p.wrap_w);
        if (sr >= 0 && sr < p.hs && sc >= 0 && sc < p.ws) val = __ldg(&a[(size_t)sr * p.ws + sc]);
      }
      S[idx] = val;
    }
    __syncthreads();
    double* d = p.dst + (size_t)img * dst_sz;
    for (int task = threadIdx.x; task < th * groups; task += blockDim.x) {
      const int g = task / th, i = task - g * th;   // adjacent lanes -> adjacent rows (bank-conflict free)
      const int j0 = g * TJ;
      double acc[TJ];
#pragma unroll
      for (int t = 0; t < TJ; ++t) acc[t] = 0.0;
      for (int kr = 0; kr < p.hk; ++kr) {
        const double* row = S + (size_t)(i + kr) * sp_w + j0;
        const double* kv = kflip + (size_t)kr * wk_pad;
        double w[2 * TJ];
#pragma unroll
        for (int t = 0; t < TJ; t += 2) {
          double2 x = *reinterpret_cast<const double2*>(row + t);
          w[t] = x.x; w[t + 1] = x.y;
        }
        for (int l0 = 0; l0 < wk_pad; l0 += TJ) {
#pragma unroll
          for (int t = 0; t < TJ; t += 2) {
            double2 x = *reinterpret_cast<const double2*>(row + l0 + TJ + t);
            w[TJ + t] = x.x; w[TJ + t + 1] = x.y;
          }
#pragma unroll
          for (int dl = 0; dl < TJ; ++dl) {
            const double kk = kv[l0 + dl];
#pragma unroll
            for (int t = 0; t < TJ; ++t) acc[t] = fma(kk, w[dl + t], acc[t]);
          }
#pragma unroll
          for (int t = 0; t < TJ; ++t) w[t] = w[TJ + t];
        }
      }
      double* o = d + (size_t)(i0 + i) * p.wd + j0;
#pragma unroll
      for (int t = 0; t < TJ; ++t)
        if (j0 + t < p.wd) o[t] = acc[t];
    }
  }
}

// Shared memory the direct kernel needs for a given row tile; sp_w is returned through *sp_w_out.
size_t direct_smem_bytes(int hk, int wk, int wd, int tile_h, int* sp_w_out) {
  const int wk_pad = (wk + TJ - 1) / TJ * TJ;
  int sp_w = (wd + TJ - 1) / TJ * TJ + wk_pad + TJ;
  while (sp_w % 16 != 2) ++sp_w;  // row stride == 4 (mod 32) words: 8 lanes x 128-bit hit 32 banks
  if (sp_w_out) *sp_w_out = sp_w;
  return sizeof(double) * ((size_t)hk * wk_pad + (size_t)(tile_h + hk - 1) * sp_w);
}

cudaError_t launch_direct(const DirectParams& p, int grid, size_t smem, cudaStream_t st) {
  cudaFuncSetAttribute(direct_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  direct_kernel<<<grid, 256, smem, st>>>(p);
  return cudaGetLastError();
}

int occupancy_direct(size_t smem) {
  int n = 0;
  cudaFuncSetAttribute(direct_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, direct_kernel, 256, smem);
  return n;
}

}  // namespace b2c
