// Microbenchmark of the tuned kernel's phase structure: per task a warp does NL x LDS.128, a block of
// NF dependent-ish FP64 instructions (ILP 12 like a radix-12 butterfly), NL x STS.128; optionally a
// __syncthreads() every SYNC tasks.  2 CTAs/SM x 8 warps (like conv_tuned108) or other shapes.
#include <cstdio>
#include <cuda_runtime.h>

template <int NL, int NFROUNDS, int SYNC>
__global__ void __launch_bounds__(256, 2) k(double* out, int tasks, long long* cyc) {
  extern __shared__ double2 sm[];
  for (int i = threadIdx.x; i < 6144; i += blockDim.x) sm[i] = make_double2(i * 1e-3, -i * 1e-3);
  __syncthreads();
  long long t0 = clock64();
  int base = threadIdx.x;
  for (int it = 0; it < tasks; ++it) {
    double2 v[NL];
#pragma unroll
    for (int q = 0; q < NL; ++q) v[q] = sm[(base + q * 256) % 6144];
#pragma unroll
    for (int r = 0; r < NFROUNDS; ++r) {
#pragma unroll
      for (int q = 0; q < NL; ++q) {   // 2 FP64 per element per round, dependent across rounds
        const double2 a = v[q], b = v[(q + 1) % NL];
        v[q].x = fma(a.x, 0.999, b.y);
        v[q].y = fma(a.y, 1.001, -b.x);
      }
    }
#pragma unroll
    for (int q = 0; q < NL; ++q) sm[(base + q * 256) % 6144] = v[q];
    base = (base + 37) % 256 + (threadIdx.x & ~255);
    if (SYNC > 0 && (it % SYNC) == SYNC - 1) __syncthreads();
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = sm[threadIdx.x].x;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int NL, int NFROUNDS, int SYNC>
void run(int tasks) {
  double* out; long long* cyc; long long h;
  cudaMalloc(&out, sizeof(double) * 296 * 256); cudaMalloc(&cyc, 8);
  auto kern = k<NL, NFROUNDS, SYNC>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 6144 * 16);
  kern<<<296, 256, 6144 * 16>>>(out, 10, cyc);
  cudaDeviceSynchronize();
  kern<<<296, 256, 6144 * 16>>>(out, tasks, cyc);
  cudaDeviceSynchronize();
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  // per SM: 16 warps x tasks; l1tex wavefronts per task = 2*NL*4 ; fp64 instr per task = NL*2*NFROUNDS
  const double wl = 16.0 * tasks * 2 * NL * 4, fp = 16.0 * tasks * NL * 2 * NFROUNDS * 0.5;
  printf("NL=%2d FP64/task=%3d sync=%d : %9lld cyc | l1tex-bound %8.0f (%.0f%%)  fp64-bound %8.0f (%.0f%%)  sum %.0f%%\n", NL,
         NL * 2 * NFROUNDS, SYNC, h, wl, 100 * wl / h, fp, 100 * fp / h, 100 * (wl + fp) / h);
  cudaFree(out); cudaFree(cyc);
}

int main() {
  run<12, 6, 0>(600); run<12, 6, 2>(600); run<12, 6, 1>(600);
  run<9, 6, 0>(600); run<9, 6, 3>(600);
  run<12, 3, 0>(600); run<12, 3, 2>(600);
  run<12, 10, 0>(600); run<12, 10, 2>(600);
  return 0;
}
