// Headline kernel: fused convolution at period 108 x 108 with a shared kernel (BASELINE configs[1]:
// 100 x 100 (*) 15 x 15 LINEAR_SAME needs a period >= 107).  Five in-place passes over one shared buffer, prime-factor
// 4 x 27 transforms without inter-stage twiddles -- the passes themselves are in conv_pfa108.cuh.
// One persistent CTA of 224 threads per image slot, two CTAs per SM.
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

// ---- the kernel spectrum as a per-thread constant cache in tensor memory ---------------------------------------
// A thread of pass 3 multiplies by the SAME 27 spectrum values for every image of the launch, but a 27-point
// transform of complex doubles already fills 108 of its 128 registers, so at most three of those values can be in
// flight from L2 at a time (96.8 KB of spectrum do not stay in the 60 KB of L1 beside two images' streams): measured,
// the exposed L2 latencies were the kernel's top stall.  Tensor memory (256 KB per SM, idle here: no tensor-core math
// in this engine) holds them instead: 108 columns per thread, written once per CTA, read back with 12-cycle latency
// and no L1/L2 traffic at all.  Two CTAs of 256 columns fill the SM's 512 columns exactly.
constexpr int TMEM_COLS = 256;
#ifndef B2C_PFA_TMEM
#define B2C_PFA_TMEM 0   // 1: spectrum values from tensor memory (see above; measured slower: only one CTA stays resident)
#endif
#ifndef B2C_PFA_MINB
#define B2C_PFA_MINB 2   // CTAs per SM the register allocation aims at
#endif

__device__ __forceinline__ unsigned tmem_alloc_cta(unsigned* slot_smem, int t) {
  if (t < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(slot_smem)), "n"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  return *slot_smem;
}
__device__ __forceinline__ void tmem_free_cta(unsigned base, int t) {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (t < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(base), "n"(TMEM_COLS) : "memory");
}
// one complex double = four 32-bit columns of the thread's own tensor-memory lane
__device__ __forceinline__ void tmem_store(unsigned addr, cplx w) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(__double2loint(w.x)),
               "r"(__double2hiint(w.x)), "r"(__double2loint(w.y)), "r"(__double2hiint(w.y)) : "memory");
}
// three complex doubles (12 columns) in one go, then wait for them
__device__ __forceinline__ void tmem_load3(unsigned addr, cplx& a, cplx& b, cplx& c) {
  unsigned r[12];
  asm volatile(
      "{\n .reg .b32 a1, a2;\n add.u32 a1, %12, 4;\n add.u32 a2, %12, 8;\n"
      " tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%12];\n"
      " tcgen05.ld.sync.aligned.32x32b.x4.b32 {%4, %5, %6, %7}, [a1];\n"
      " tcgen05.ld.sync.aligned.32x32b.x4.b32 {%8, %9, %10, %11}, [a2];\n"
      " tcgen05.wait::ld.sync.aligned;\n}"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11])
      : "r"(addr));
  a = make_double2(__hiloint2double((int)r[1], (int)r[0]), __hiloint2double((int)r[3], (int)r[2]));
  b = make_double2(__hiloint2double((int)r[5], (int)r[4]), __hiloint2double((int)r[7], (int)r[6]));
  c = make_double2(__hiloint2double((int)r[9], (int)r[8]), __hiloint2double((int)r[11], (int)r[10]));
}
// this thread's tensor-memory address: lanes 32 (w % 4).. belong to warp w; warps 4..6 use the columns after those of
// warps 0..3.  Re-derived from the slot in shared memory where it is needed: nothing stays in registers.
__device__ __forceinline__ unsigned tmem_addr(const unsigned* slot_smem, int t) {
  return *(const volatile unsigned*)slot_smem + ((unsigned)((t >> 5) & 3) << 21) + (unsigned)(t >> 7) * 108u;
}
// v *= kernel spectrum (swapped for the inverse transform), spectrum from tensor memory, three values at a time
__device__ __forceinline__ void p3_mul_tmem(cplx* v, unsigned taddr) {
#pragma unroll
  for (int g = 0; g < 27; g += 3) {
    cplx a, b, c;
    tmem_load3(taddr + 4 * g, a, b, c);
    v[g] = cswap(cmul(v[g], a));
    v[g + 1] = cswap(cmul(v[g + 1], b));
    v[g + 2] = cswap(cmul(v[g + 2], c));
  }
}

// The four 27-point transforms a thread runs per image (pass 1, pass 3 forward, pass 3 backward, pass 5) go
// through ONE copy of dft27 inside a four-step loop: what differs per step is only where the 27 values come from
// and go to.  Inlined four times the kernel is 140 KB of code and the warps of two CTAs, each in another pass,
// starve on instruction fetch (measured: stall_no_instruction 2.1 cycles per issue, 13.1 M conv/s).
template <bool EXT>
__global__ void __launch_bounds__(THREADS, B2C_PFA_MINB) conv_kernel(const __grid_constant__ Params p) {
  cplx* Z = reinterpret_cast<cplx*>(b2c_smem_pfa);
  const int t = threadIdx.x;
  const VThread q = vthread(t);
  cplx* z3 = p3_base(t, Z);
  // this thread's 27 spectrum values -> its tensor-memory lane (lanes 32 (w % 4).. of warp w; warps 4..6 use the
  // columns after those of warps 0..3)
#if B2C_PFA_TMEM
  __shared__ unsigned tmem_slot;
  tmem_alloc_cta(&tmem_slot, t);
  const unsigned taddr = tmem_addr(&tmem_slot, t);
  {
    const cplx* __restrict__ ks = p.kspec + t / KK * (27 * KK) + t % KK;
#pragma unroll 1
    for (int k = 0; k < 27; ++k) tmem_store(taddr + 4 * k, __ldg(ks + k * KK));
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
#endif
  const size_t src_sz = (size_t)p.hs * p.ws, dst_sz = (size_t)p.hd * p.wd;
  for (long img = blockIdx.x; img < p.n_images; img += gridDim.x) {
#ifdef B2C_PFA_TRACE
    const bool traced = img == blockIdx.x + 3L * gridDim.x && blockIdx.x < 512;
#endif
    B2C_STAMP(0)
    if (t == 0 && img + gridDim.x < p.n_images)   // next image -> L2
      l2_prefetch(p.src + (size_t)(img + gridDim.x) * src_sz, src_sz * sizeof(double));
#pragma unroll 1
    for (int step = 0; step < 3; ++step) {   // pass 1, pass 3, pass 5
      if (step == 1 || q.act) {
        cplx v[27];   // defined by every prologue, consumed by every epilogue: not live across passes 2 and 4
        if (step == 0) p1_load<EXT>(v, q, p, p.src + (size_t)img * src_sz);
        else if (step == 1) p3_load(v, z3);
        else p5_load(v, q, Z);
#pragma unroll 1
        for (int rep = (step == 1 ? 0 : 1); rep < 2; ++rep) {   // pass 3 transforms twice: forward, spectrum, backward
          #if B2C_PFA_TMEM
          if (step == 1 && rep == 1) p3_mul_tmem(v, taddr);
#else
          if (step == 1 && rep == 1) p3_mul(v, t, p.kspec);
#endif
          dft27(v);
        }
        if (step == 0) p1_store(v, q, Z);
        else if (step == 1) p3_store(v, z3);
        else p5_store(v, q, p, p.dst + (size_t)img * dst_sz);
      }
      if (step == 0) {
        B2C_STAMP(1)
        __syncthreads();
        B2C_STAMP(2)
        pass2(t, Z);
        B2C_STAMP(3)
        __syncthreads();
        B2C_STAMP(4)
      } else if (step == 1) {
        B2C_STAMP(5)
        __syncthreads();
        B2C_STAMP(6)
        pass4(t, Z);
        B2C_STAMP(7)
        __syncthreads();
        B2C_STAMP(8)
      }
    }
    B2C_STAMP(9)
    // no barrier: the next image's pass 1 writes exactly the slots this thread just read
  }
#if B2C_PFA_TMEM
  tmem_free_cta(tmem_slot, t);
#endif
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
  cudaFuncSetAttribute(pfa::conv_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pfa::SMEM_BYTES);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, pfa::conv_kernel<false>, pfa::THREADS, pfa::SMEM_BYTES);
#ifdef B2C_PFA_FORCE_OCC
  occ = B2C_PFA_FORCE_OCC;   // A/B measurements only
#endif
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
  if (sh_h | sh_w) {
    cudaFuncSetAttribute(pfa::conv_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pfa::SMEM_BYTES);
    pfa::conv_kernel<true><<<grid, pfa::THREADS, pfa::SMEM_BYTES, st>>>(p);
  } else {
    cudaFuncSetAttribute(pfa::conv_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pfa::SMEM_BYTES);
    pfa::conv_kernel<false><<<grid, pfa::THREADS, pfa::SMEM_BYTES, st>>>(p);
  }
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
