// Direct (nested-loop) convolution kernel for small kernels, where the reference itself finds
// loops faster than the FFT (README.md:31,47; Convolution/src/convolution_std.h:86-199 is the
// CPU form).
//
// One CTA stages a row tile of the source (zero- or wrap-padded halo) and the flipped kernel in
// shared memory.  A thread owns ONE output column and TI = 8 consecutive output rows; lanes of a
// warp walk adjacent columns, so shared-memory reads are conflict-free 64-bit accesses and global
// stores are coalesced.  For every kernel column the thread loads a vertical window of
// TI + HK - 1 source values into registers and performs TI * HK FMAs on it; HK (the kernel height)
// is a template parameter so that the window lives in registers.  Kernels taller than
// kMaxDirectHK use the general variant (one load per FMA), which only matters for forced-path
// tests: the planner never picks the direct path there.
#include "conv_direct_bulk.cuh"
#include "launch_cache.h"
#include "conv_params.h"

namespace b2c {

// HK > 0: compile-time kernel height; HK == 0: run-time height (general variant).
// TI: output rows per thread (8 for images, 1 for 1-D row signals and very short images).
template <int HK, int TI, bool WRAP>
__global__ void __launch_bounds__(kDirectThreads, 2) direct_kernel(const __grid_constant__ DirectParams p) {
  double* kflip = reinterpret_cast<double*>(b2c_smem_direct);   // [wk][hk_pad]: column-major, flipped
  const int hk = HK > 0 ? HK : p.hk;
  const int hk_pad = (hk + 1) & ~1;
  double* S = kflip + (size_t)p.wk * hk_pad;
  const int pitch = p.sp_w;
  const size_t src_sz = (size_t)p.hs * p.ws, ker_sz = (size_t)p.hk * p.wk, dst_sz = (size_t)p.hd * p.wd;
  // A work unit is one row tile of one image, or -- for small images with a shared kernel -- the whole
  // of `group` consecutive images (their pixels are contiguous in memory: coalesced staging and stores,
  // and enough tasks for all 256 threads).
  const int group = p.group;
  const long units = group > 1 ? (p.n_images + group - 1) / group : p.n_images * p.tiles_per_image;
  bool kernel_loaded = false;

  for (long work = blockIdx.x; work < units; work += gridDim.x) {
    const long img0 = group > 1 ? work * group : work / p.tiles_per_image;
    const int ng = group > 1 ? (int)min((long)group, p.n_images - img0) : 1;
    const int tile = group > 1 ? 0 : (int)(work - img0 * p.tiles_per_image);
    const int i0 = tile * p.tile_h;
    const int th = min(p.tile_h, p.hd - i0);
    __syncthreads();  // previous iteration's readers are done
    if (!(p.kernel_is_shared && kernel_loaded)) {
      const double* k = p.ker + (p.kernel_is_shared ? 0 : (size_t)img0 * ker_sz);
      for (int idx = threadIdx.x; idx < p.wk * hk_pad; idx += kDirectThreads) {
        const int c = idx / hk_pad, r = idx - c * hk_pad;
        kflip[idx] = r < hk ? __ldg(&k[(size_t)(hk - 1 - r) * p.wk + (p.wk - 1 - c)]) : 0.0;
      }
      kernel_loaded = true;
    }
    // stage source rows [i0 + off_h - (hk-1), ...) with halo.  Eight independent global loads are in
    // flight per thread before the first shared store (the staging loop is latency-, not bandwidth-bound).
    const double* a = p.src + (size_t)img0 * src_sz;
    const int rows = th + hk - 1;
    const int cols = p.wd + p.wk - 1;
    const int nelem = rows * pitch;            // staged doubles per image
    const int nall = ng * nelem;
    constexpr int SU = 8;   // independent global loads in flight per thread before the first shared store
    const int sr0 = i0 + p.off_h - (hk - 1), sc0 = p.off_w - (p.wk - 1);
    const int step_u = kDirectThreads / pitch, step_v = kDirectThreads - step_u * pitch;   // 256 = step_u * pitch + step_v
    if (group > 1 && 2 * step_u > rows) {
      // several very small images: flat index -> (image, row, column) by exact float-reciprocal divisions
      const float inv_pitch = 1.0f / (float)pitch, inv_nelem = 1.0f / (float)nelem;
      for (int base = threadIdx.x; base < nall; base += SU * kDirectThreads) {
        double vals[SU];
#pragma unroll
        for (int k = 0; k < SU; ++k) {
          const int idx = base + k * kDirectThreads;
          double val = 0.0;
          if (idx < nall) {
            const int g = fast_div(idx, nelem, inv_nelem);
            const int rem = idx - g * nelem;
            const int u = fast_div(rem, pitch, inv_pitch), v = rem - u * pitch;
            int sr = sr0 + u, sc = sc0 + v;
            if (WRAP) { sr = wrap_index(sr, p.wrap_h); sc = wrap_index(sc, p.wrap_w); }
            if (v < cols && (unsigned)sr < (unsigned)p.hs && (unsigned)sc < (unsigned)p.ws)
              val = __ldg(a + ((long)g * (long)src_sz + (long)(sr * p.ws + sc)));
          }
          vals[k] = val;
        }
#pragma unroll
        for (int k = 0; k < SU; ++k) {
          const int idx = base + k * kDirectThreads;
          if (idx < nall) S[idx] = vals[k];
        }
      }
    } else {
      // each thread walks idx = tid, tid + 256, ... keeping (image g, row u, column v) by increments
      // (no division per element: the flat-index form made the kernel issue-bound)
      int u = threadIdx.x / pitch, v = threadIdx.x - u * pitch;
      const double* ag = a;
      if (group > 1) { const int g0 = u / rows; u -= g0 * rows; ag += (size_t)g0 * src_sz; }
      // (one 8-byte cp.async per element, zero-filled outside the source: no registers, nothing to wait for until the
      // whole tile has been requested; measured against eight register-staged loads per thread: 99 x 99 (*) 3 x 3
      // 20.7 -> 23.6 M conv/s, CIRCULAR_SAME 100 x 100 (*) 3 x 3 13.6 -> 17.2 M, tools/bench_direct_odd.py)
      for (int idx = threadIdx.x; idx < nall; idx += kDirectThreads) {
        int sr = sr0 + u, sc = sc0 + v;
        if (WRAP) { sr = wrap_index(sr, p.wrap_h); sc = wrap_index(sc, p.wrap_w); }
        const bool in = v < cols && (unsigned)sr < (unsigned)p.hs && (unsigned)sc < (unsigned)p.ws;
        cp_async8(S + idx, in ? ag + (sr * p.ws + sc) : a, in ? 8 : 0);
        v += step_v; u += step_u;
        if (v >= pitch) { v -= pitch; ++u; }
        if (group > 1 && u >= rows) { u -= rows; ag += src_sz; }   // 2 * step_u <= rows: at most one image per step
      }
      cp_async_wait_all();
    }
    __syncthreads();
    double* d = p.dst + (size_t)img0 * dst_sz;
    const int blocks = (th + TI - 1) / TI;
    const int per_image = blocks * p.wd;
    const float inv_per_image = 1.0f / (float)per_image, inv_wd = 1.0f / (float)p.wd;
    for (int task = threadIdx.x; task < ng * per_image; task += kDirectThreads) {
      const int g = group > 1 ? fast_div(task, per_image, inv_per_image) : 0;
      const int t2 = task - g * per_image;
      const int b = fast_div(t2, p.wd, inv_wd), j = t2 - b * p.wd;   // adjacent lanes -> adjacent columns
      const int r0 = b * TI;
      double acc[TI];
#pragma unroll
      for (int t = 0; t < TI; ++t) acc[t] = 0.0;
      const double* col = S + (g * nelem + r0 * pitch + j);       // 32-bit offsets into shared memory
      if (HK > 0) {
        const int last = rows - 1 - r0;   // highest staged row of this window (rows past it feed no existing output)
        // (a clamp-free copy of this loop for windows that are fully staged measured slower: lanes of a warp
        // straddle both cases and run both)
        for (int kc = 0; kc < p.wk; ++kc) {
          double w[TI + (HK > 0 ? HK : 1) - 1];
#pragma unroll
          for (int u = 0; u < TI + HK - 1; ++u) w[u] = col[min(u, last) * pitch + kc];
          const double* kv = kflip + kc * hk_pad;
#pragma unroll
          for (int kr = 0; kr < HK; ++kr) {
            const double kk = kv[kr];
#pragma unroll
            for (int t = 0; t < TI; ++t) acc[t] = fma(kk, w[t + kr], acc[t]);
          }
        }
      } else {
        const int nt = min(TI, th - r0);
        for (int kc = 0; kc < p.wk; ++kc) {
          const double* kv = kflip + kc * hk_pad;
          for (int kr = 0; kr < hk; ++kr) {
            const double kk = kv[kr];
#pragma unroll
            for (int t = 0; t < TI; ++t)
              if (t < nt) acc[t] = fma(kk, col[(t + kr) * pitch + kc], acc[t]);
          }
        }
      }
      double* o = d + (size_t)g * dst_sz + (size_t)(i0 + r0) * p.wd + j;
#pragma unroll
      for (int t = 0; t < TI; ++t)
        if (r0 + t < th) o[(size_t)t * p.wd] = acc[t];
    }
  }
}

// Can the bulk-staged variant run this geometry?  Fills the staged pitch, the column shift and the bytes of shared
// memory.  Conditions: every row copy starts and ends on a 16-byte boundary in global AND shared memory.
// Output rows per thread of the bulk-staged kernel: the member of {8, 10, 12} that wastes the fewest thread-rows on a
// work unit -- rows past the tile in its last block plus idle threads in the last round of tasks.  (100 rows: 13
// blocks of 8 are 104 rows and 1300 tasks = 5.08 rounds of 256 threads, 81 % useful; 10 blocks of 10 are 1000 tasks =
// 3.9 rounds, 98 %.  Measured at 100 x 100 (*) 7 x 7: 14.6 -> 15.7 M conv/s; 96 rows want 12, 64 rows 8.)
int direct_bulk_ti(int hd, int wd, int tile_h, int group) {
  if (hd < 8) return 1;
  const int th = tile_h < hd ? tile_h : hd;
  int best = 8;
  double best_eff = 0.0;
  for (int ti : {8, 10, 12}) {
    const long tasks = (long)((th + ti - 1) / ti) * wd * group;
    const double eff = (double)th * wd * group / ((double)((tasks + kDirectThreads - 1) / kDirectThreads) * kDirectThreads * ti);
    if (eff > best_eff + 1e-9) { best_eff = eff; best = ti; }
  }
  return best;
}

// Geometry of the odd-width form (direct_bulk_kernel<..., ODD = true>): whole rows, one spare staged column for the
// per-row shift.
bool direct_bulk_geometry_odd(int hs, int ws, int hk, int wk, int hd, int wd, int off_h, int off_w, int tile_h, int group,
                              int nbuf, int ti, int* pitch, size_t* smem) {
  (void)hs; (void)hd; (void)off_h;
  if (hk < 1 || hk > kMaxDirectHK || !(ws & 1) || ws < 3) return false;
  const int sc0 = off_w - (wk - 1), cols = wd + wk - 1;
  if (sc0 > 0 || sc0 + cols < ws) return false;                // whole rows are staged
  const int hk_pad = (hk + 1) & ~1;
  if ((wk * hk_pad) & 1) return false;                         // S itself starts 16-byte aligned behind kflip
  const int rows = (tile_h + ti - 1) / ti * ti + hk - 1;
  const int pt = (cols + 1 + 1) & ~1;
  *pitch = pt;
  *smem = 16 + sizeof(double) * ((size_t)wk * hk_pad + (size_t)nbuf * group * rows * pt);
  return true;
}

bool direct_bulk_geometry(int hs, int ws, int hk, int wk, int hd, int wd, int off_h, int off_w, int tile_h, int group,
                          int nbuf, int ti, int* pitch, int* col_shift, size_t* smem) {
  if (hk < 1 || hk > kMaxDirectHK || (ws & 1)) return false;
  if (group > 1 && (((size_t)hs * ws) & 1)) return false;     // every image of a group starts 16-byte aligned
  const int sc0 = off_w - (wk - 1), cols = wd + wk - 1;
  const int c_lo = sc0 > 0 ? sc0 : 0, c_hi = ws < sc0 + cols ? ws : sc0 + cols;
  const int rows = (tile_h + ti - 1) / ti * ti + hk - 1;   // as direct_bulk_kernel pads them
  if (c_hi <= c_lo) return false;
  if ((c_lo & 1) || ((c_hi - c_lo) & 1)) return false;        // source side: ws is even, so rows start aligned
  const int shift = (c_lo - sc0) & 1;                          // destination side: staged column c_lo - sc0 + shift is even
  const int hk_pad = (hk + 1) & ~1;
  if ((wk * hk_pad) & 1) return false;                         // S itself starts 16-byte aligned behind kflip
  int pt = cols + shift;
  pt = (pt + 1) & ~1;
  *pitch = pt; *col_shift = shift;
  *smem = 16 + sizeof(double) * ((size_t)wk * hk_pad + (size_t)nbuf * group * rows * pt);
  return true;
}

// Shared memory the direct kernel needs for a given row tile; the staged row pitch is returned
// through *sp_w_out.
size_t direct_smem_bytes(int hk, int wk, int wd, int tile_h, int* sp_w_out, int group) {
  const int hk_pad = (hk + 1) & ~1;
  int pitch = wd + wk - 1;
  pitch = (pitch + 1) & ~1;
  if (sp_w_out) *sp_w_out = pitch;
  return sizeof(double) * ((size_t)wk * hk_pad + (size_t)group * (tile_h + hk - 1) * pitch);
}


template <int HK>
static direct_fn pick(int ti, bool wrap) {
  if (HK == 0) ti = ti > 1 ? 8 : 1;   // run-time kernel height: two sizes only
  switch (ti) {
    case 10: return wrap ? direct_kernel<HK, (HK ? 10 : 8), true> : direct_kernel<HK, (HK ? 10 : 8), false>;
    case 12: return wrap ? direct_kernel<HK, (HK ? 12 : 8), true> : direct_kernel<HK, (HK ? 12 : 8), false>;
    case 1: return wrap ? direct_kernel<HK, 1, true> : direct_kernel<HK, 1, false>;
  }
  return wrap ? direct_kernel<HK, 8, true> : direct_kernel<HK, 8, false>;
}

// ti: output rows per thread (direct_bulk_ti: 1, 8, 10 or 12)
static direct_fn direct_variant(int hk, int ti, bool wrap) {
  switch (hk) {
    case 1: return pick<1>(ti, wrap);   case 2: return pick<2>(ti, wrap);   case 3: return pick<3>(ti, wrap);   case 4: return pick<4>(ti, wrap);
    case 5: return pick<5>(ti, wrap);   case 6: return pick<6>(ti, wrap);   case 7: return pick<7>(ti, wrap);   case 8: return pick<8>(ti, wrap);
    case 9: return pick<9>(ti, wrap);   case 10: return pick<10>(ti, wrap); case 11: return pick<11>(ti, wrap); case 12: return pick<12>(ti, wrap);
    case 13: return pick<13>(ti, wrap); case 14: return pick<14>(ti, wrap); case 15: return pick<15>(ti, wrap); case 16: return pick<16>(ti, wrap);
  }
  return pick<0>(ti, wrap);
}

// ti: output rows per thread (direct_bulk_ti); odd: the odd-width form
direct_fn direct_bulk_variant_odd(int hk, int nbuf, int ti);   // conv_direct_odd.cu
static direct_fn direct_bulk_variant(int hk, int nbuf, int ti, bool odd) {
  return odd ? direct_bulk_variant_odd(hk, nbuf, ti) : direct_bulk_variant_o<false>(hk, nbuf, ti);
}

int occupancy_direct_bulk(int hk, int ti, int nbuf, size_t smem, bool odd) {
  direct_fn fn = direct_bulk_variant(hk, nbuf, ti, odd);
  if (!fn) return 0;
  return cached_occupancy(reinterpret_cast<const void*>(fn), kDirectThreads, smem);
}

cudaError_t launch_direct(const DirectParams& p, int grid, size_t smem, cudaStream_t st) {
  if (p.bulk) {
    direct_fn fn = direct_bulk_variant(p.hk, p.bulk, p.ti, p.odd != 0);
    ensure_dyn_smem(reinterpret_cast<const void*>(fn), smem);
    fn<<<grid, kDirectThreads, smem, st>>>(p);
    return cudaGetLastError();
  }
  direct_fn fn = direct_variant(p.hk, p.ti, p.wrap_h > 0);
  ensure_dyn_smem(reinterpret_cast<const void*>(fn), smem);
  fn<<<grid, kDirectThreads, smem, st>>>(p);
  return cudaGetLastError();
}

int occupancy_direct(int hk, int ti, bool wrap, size_t smem) {
  int n = 0;
  direct_fn fn = direct_variant(hk, ti, wrap);
  n = cached_occupancy(reinterpret_cast<const void*>(fn), kDirectThreads, smem);
  return n;
}

}  // namespace b2c
