// Microbenchmark: FP64 FMA throughput vs ILP / warps per SM, and overlap of FP64 with 128-bit
// shared-memory traffic on B200.  Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o fp64_lds fp64_lds.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int ILP, bool DO_FMA, bool DO_LDS>
__global__ void k(double* out, int iters, long long* cyc) {
  extern __shared__ double2 sm[];
  double a[ILP > 0 ? ILP : 1];
  double2 acc = make_double2(0, 0);
#pragma unroll
  for (int i = 0; i < ILP; ++i) a[i] = threadIdx.x * 1e-3 + i;
  const double b = 1.0000001, c = 1e-9;
  for (int i = threadIdx.x; i < 2048; i += blockDim.x) sm[i] = make_double2(i, -i);
  __syncthreads();
  long long t0 = clock64();
  int idx = threadIdx.x;
  for (int it = 0; it < iters; ++it) {
    if (DO_FMA) {
#pragma unroll
      for (int r = 0; r < 8; ++r)
#pragma unroll
        for (int i = 0; i < ILP; ++i) a[i] = fma(a[i], b, c);
    }
    if (DO_LDS) {
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        double2 v = sm[(idx + r * 256) & 2047];
        acc.x += v.x; acc.y += v.y;   // 2 DADD per LDS (kept small)
      }
      idx += 32;
    }
  }
  long long t1 = clock64();
  double s = acc.x + acc.y;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int ILP, bool F, bool L>
void run(const char* name, int warps_per_sm, int iters) {
  double* out; long long* cyc; long long h;
  int threads = 32 * warps_per_sm;   // one CTA per SM
  cudaMalloc(&out, sizeof(double) * 148 * threads); cudaMalloc(&cyc, 8);
  k<ILP, F, L><<<148, threads, 2048 * 16>>>(out, 10, cyc);
  cudaDeviceSynchronize();
  k<ILP, F, L><<<148, threads, 2048 * 16>>>(out, iters, cyc);
  cudaDeviceSynchronize();
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  double fma_per_clk_sm = F ? (double)iters * 8 * ILP * threads / h : 0;
  double lds_wave_per_clk = L ? (double)iters * 4 * warps_per_sm * 4 / h : 0;   // 4 wavefronts per LDS.128
  printf("%-10s ILP=%d warps/SM=%2d : %9lld cyc  FMA/clk/SM=%6.1f  LDS wavefronts/clk=%5.2f\n", name, ILP, warps_per_sm, h,
         fma_per_clk_sm, lds_wave_per_clk);
  cudaFree(out); cudaFree(cyc);
}

int main() {
  const int it = 2000;
  for (int w : {4, 8, 16, 32}) {
    run<1, true, false>("fma", w, it); run<2, true, false>("fma", w, it); run<4, true, false>("fma", w, it);
    run<8, true, false>("fma", w, it);
  }
  for (int w : {4, 8, 16, 32}) run<1, false, true>("lds", w, it);
  for (int w : {8, 16, 32}) { run<4, true, true>("fma+lds", w, it); run<8, true, true>("fma+lds", w, it); }
  return 0;
}
