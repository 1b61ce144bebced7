// Microbenchmark: 128-bit shared-memory load vs store throughput (wavefronts per clock per SM) on B200.
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>   // 0: LDS.128 only, 1: STS.128 only, 2: LDS+STS alternating (in-place update), 3: STS.64 only, 4: LDS.64 only
__global__ void k(double* out, int iters, long long* cyc) {
  extern __shared__ double2 sm[];
  for (int i = threadIdx.x; i < 2048; i += blockDim.x) sm[i] = make_double2(i, -i);
  __syncthreads();
  double2 acc = make_double2(threadIdx.x, 1.0);
  double* smd = reinterpret_cast<double*>(sm);
  int idx = threadIdx.x;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const int a = (idx + r * 256) & 2047;
      if (MODE == 0) { double2 v = sm[a]; acc.x += v.x; }
      if (MODE == 1) { sm[a] = acc; }
      if (MODE == 2) { double2 v = sm[a]; v.x += 1.0; sm[a] = v; }
      if (MODE == 3) { smd[a] = acc.x; }
      if (MODE == 4) { acc.x += smd[a]; }
    }
    idx += 32;
    if (MODE == 1 || MODE == 3) acc.x += 1.0;
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc.x + acc.y + sm[threadIdx.x].x;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int MODE>
void run(const char* name, int warps, int iters, double wave_per_instr, double instr_per_r) {
  double* out; long long* cyc; long long h;
  int threads = 32 * warps;
  cudaMalloc(&out, sizeof(double) * 148 * threads); cudaMalloc(&cyc, 8);
  k<MODE><<<148, threads, 2048 * 16>>>(out, 10, cyc);
  cudaDeviceSynchronize();
  k<MODE><<<148, threads, 2048 * 16>>>(out, iters, cyc);
  cudaDeviceSynchronize();
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("%-14s warps/SM=%2d : %9lld cyc  -> %.3f cycles per wavefront (min-wavefront model)\n", name, warps, h,
         (double)h / ((double)iters * 8 * warps * wave_per_instr * instr_per_r));
  cudaFree(out); cudaFree(cyc);
}

int main() {
  const int it = 2000;
  for (int w : {8, 16, 32}) {
    run<0>("LDS.128", w, it, 4, 1);
    run<1>("STS.128", w, it, 4, 1);
    run<2>("LDS+STS.128", w, it, 4, 2);
    run<3>("STS.64", w, it, 2, 1);
    run<4>("LDS.64", w, it, 2, 1);
  }
  return 0;
}
