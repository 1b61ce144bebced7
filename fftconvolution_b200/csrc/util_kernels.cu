// Small utility kernels: counter-based synthetic data and the
// order-independent verification checksum the multi-GPU partitioner reduces.
#include <cuda_runtime.h>

#include "fft_device.cuh"

namespace b2c {

__device__ __forceinline__ unsigned long long splitmix64(unsigned long long x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

// d[i] = (splitmix64(seed + first + i) >> 11) * 2^-53  in [0,1)
__global__ void fill_uniform_kernel(double* __restrict__ d, long n, unsigned long long seed, long first) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    d[i] = (double)(splitmix64(seed + (unsigned long long)(first + i)) >> 11) * (1.0 / 9007199254740992.0);
}

__global__ void checksum_kernel(const double* __restrict__ d, long n, unsigned long long* __restrict__ out3) {
  unsigned long long s = 0, x = 0;
  double fs = 0.0;
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const double v = d[i];
    const unsigned long long b = (unsigned long long)__double_as_longlong(v);
    s += b;
    x ^= b;
    fs += v;
  }
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    x ^= __shfl_xor_sync(0xffffffffu, x, o);
    fs += __shfl_xor_sync(0xffffffffu, fs, o);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(&out3[0], s);
    atomicXor(&out3[1], x);
    atomicAdd(reinterpret_cast<double*>(&out3[2]), fs);  // informational only: order dependent
  }
}

// Per-image magnitude balance for conv_big2d.cu (see fft_device.cuh: balance_from_maxima).  The images can be
// large and few: `chunks` CTAs scan one image, maxima are combined with atomicMax on the bit patterns
// (non-negative doubles order like unsigned integers), a second tiny kernel turns them into alpha.
__global__ void __launch_bounds__(256) absmax_kernel(const double* __restrict__ src, long src_sz,
                                                     const double* __restrict__ ker, long ker_sz, long ker_stride,
                                                     unsigned long long* __restrict__ maxima) {
  __shared__ double red[2][8];
  const long img = blockIdx.y;
  const double* a = src + img * src_sz;
  const double* k = ker + img * ker_stride;
  const long stride = (long)gridDim.x * blockDim.x, first = (long)blockIdx.x * blockDim.x + threadIdx.x;
  double ms = 0.0, mk = 0.0;
  for (long i = first; i < src_sz; i += stride) ms = fmax(ms, fabs(__ldg(a + i)));
  for (long i = first; i < ker_sz; i += stride) mk = fmax(mk, fabs(__ldg(k + i)));
  for (int o = 16; o > 0; o >>= 1) {
    ms = fmax(ms, __shfl_xor_sync(0xffffffffu, ms, o));
    mk = fmax(mk, __shfl_xor_sync(0xffffffffu, mk, o));
  }
  if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = ms; red[1][threadIdx.x >> 5] = mk; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) { ms = fmax(ms, red[0][w]); mk = fmax(mk, red[1][w]); }
    if (!(ms == ms)) ms = 0.0;   // NaN input: no scaling (the result is NaN either way)
    if (!(mk == mk)) mk = 0.0;
    atomicMax(maxima + 2 * img, (unsigned long long)__double_as_longlong(ms));
    atomicMax(maxima + 2 * img + 1, (unsigned long long)__double_as_longlong(mk));
  }
}

__global__ void balance_finalize_kernel(const unsigned long long* __restrict__ maxima, long n, double* __restrict__ alpha) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n)
    alpha[i] = balance_from_maxima(__longlong_as_double((long long)maxima[2 * i]), __longlong_as_double((long long)maxima[2 * i + 1]));
}

// d_alpha: n doubles; d_maxima: 2 n words of scratch
cudaError_t launch_balance(const double* d_src, long src_sz, const double* d_ker, long ker_sz, long ker_stride, long n,
                           double* d_alpha, unsigned long long* d_maxima, int sms, cudaStream_t st) {
  cudaError_t e = cudaMemsetAsync(d_maxima, 0, sizeof(unsigned long long) * 2 * (size_t)n, st);
  if (e != cudaSuccess) return e;
  long chunks = (2L * sms + n - 1) / n;                       // about two waves of CTAs
  const long most = (src_sz + 256 * 16 - 1) / (256 * 16);     // at least 16 elements per thread
  if (chunks > most) chunks = most;
  if (chunks < 1) chunks = 1;
  for (long i0 = 0; i0 < n; i0 += 65535) {                    // gridDim.y limit
    const long ni = n - i0 < 65535 ? n - i0 : 65535;
    absmax_kernel<<<dim3((unsigned)chunks, (unsigned)ni), 256, 0, st>>>(d_src + i0 * src_sz, src_sz, d_ker + i0 * ker_stride, ker_sz,
                                                                  ker_stride, d_maxima + 2 * i0);
  }
  balance_finalize_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(d_maxima, n, d_alpha);
  return cudaGetLastError();
}

cudaError_t launch_fill_uniform(double* d, long n, unsigned long long seed, long first, cudaStream_t st) {
  int blocks = (int)((n + 255) / 256 < 148L * 16 ? (n + 255) / 256 : 148L * 16);
  fill_uniform_kernel<<<blocks, 256, 0, st>>>(d, n, seed, first);
  return cudaGetLastError();
}

cudaError_t launch_checksum(const double* d, long n, unsigned long long* d_out3, cudaStream_t st) {
  int blocks = (int)((n + 255) / 256 < 148L * 8 ? (n + 255) / 256 : 148L * 8);
  checksum_kernel<<<blocks, 256, 0, st>>>(d, n, d_out3);
  return cudaGetLastError();
}

// ---- single-precision and complex inputs: thin adapters around the FP64 real engine ----------
__global__ void widen_kernel(const float* __restrict__ in, double* __restrict__ out, long n) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = (double)__ldg(in + i);
}
__global__ void narrow_kernel(const double* __restrict__ in, float* __restrict__ out, long n) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = (float)in[i];
}
// interleaved complex -> two real planes
__global__ void split_complex_kernel(const double2* __restrict__ in, double* __restrict__ re, double* __restrict__ im, long n) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const double2 v = __ldg(in + i);
    re[i] = v.x; im[i] = v.y;
  }
}
// out[i].part = a[i] + sign * b[i]   (part 0: real, 1: imaginary; the other part is left as it is)
__global__ void combine_complex_kernel(const double* __restrict__ a, const double* __restrict__ b, double sign, double* __restrict__ out,
                                       int part, long n) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[2 * i + part] = a[i] + sign * b[i];
}

static int grid_for(long n) { return (int)((n + 255) / 256 < 148L * 16 ? (n + 255) / 256 : 148L * 16); }
cudaError_t launch_widen(const float* in, double* out, long n, cudaStream_t st) {
  if (n > 0) widen_kernel<<<grid_for(n), 256, 0, st>>>(in, out, n);
  return cudaGetLastError();
}
cudaError_t launch_narrow(const double* in, float* out, long n, cudaStream_t st) {
  if (n > 0) narrow_kernel<<<grid_for(n), 256, 0, st>>>(in, out, n);
  return cudaGetLastError();
}
cudaError_t launch_split_complex(const double* in, double* re, double* im, long n, cudaStream_t st) {
  if (n > 0) split_complex_kernel<<<grid_for(n), 256, 0, st>>>(reinterpret_cast<const double2*>(in), re, im, n);
  return cudaGetLastError();
}
cudaError_t launch_combine_complex(const double* a, const double* b, double sign, double* out, int part, long n, cudaStream_t st) {
  if (n > 0) combine_complex_kernel<<<grid_for(n), 256, 0, st>>>(a, b, sign, out, part, n);
  return cudaGetLastError();
}

}  // namespace b2c
