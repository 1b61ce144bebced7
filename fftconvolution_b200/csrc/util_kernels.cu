// Small utility kernels: counter-based synthetic data and the
// order-independent verification checksum the multi-GPU partitioner reduces.
#include <cuda_runtime.h>

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

}  // namespace b2c
