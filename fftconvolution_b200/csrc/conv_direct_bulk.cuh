// The bulk-staged direct kernel (see conv_direct.cu for the kernel family and the host side).  A header because its
// 256 instantiations (kernel height x buffers x task rows x odd-width form) are compiled in two translation units
// (conv_direct.cu: even widths, conv_direct_odd.cu: odd widths) so that the build's longest file stays short.
#pragma once
#include "bulk_copy.cuh"
#include "conv_params.h"
#include "fft_device.cuh"

namespace b2c {

extern __shared__ __align__(16) unsigned char b2c_smem_direct[];

constexpr int kMaxDirectHK = 16;
constexpr int kDirectThreads = 256;

// 8-byte asynchronous copy global -> shared; bytes = 0 writes zeros (the halo) without touching global memory
__device__ __forceinline__ void cp_async8(double* dst_smem, const double* src_gmem, int bytes) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"((unsigned)__cvta_generic_to_shared(dst_smem)), "l"(src_gmem), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>   // returns when at most the N most recent groups of this thread are still in flight
__device__ __forceinline__ void cp_async_wait_group() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ int wrap_index(int x, int m) {
  x %= m;
  return x < 0 ? x + m : x;
}

typedef void (*direct_fn)(const DirectParams);

// ---- bulk-staged variant ---------------------------------------------------------------------------------------
// Whole images with a compile-time kernel height, no wrapping: the staging loop of direct_kernel (one 8-byte load,
// ~15 instructions of index arithmetic and bounds tests per source element: ncu showed 90 instructions per OUTPUT
// pixel at 100 x 100 (*) 3 x 3, issue slots 55 % busy, HBM 41 %) is replaced by one bulk asynchronous copy per source
// row (cp.async.bulk + mbarrier, issued by as many threads as there are rows): no registers, no index arithmetic, the
// zero halo is written once per CTA and never touched again because the copies only land in the interior.
// Layout of the staged tile: row u <-> source row sr0 + u, column v <-> source column sc0 + v - col_shift
// (sr0 = off_h - (hk-1), sc0 = off_w - (wk-1); col_shift in {0, 1} makes every copy's destination 16-byte aligned).
// Work units as in direct_kernel: a row tile of one image, or `group` whole small images.  With row tiles the rows
// above / below the source are cleared per unit (they move with the tile).
// The host selects this variant only when every copy is 16-byte aligned at both ends (direct_bulk_geometry).
// NBUF = 2: two staged tiles per CTA -- the copies of the next work unit are in flight while this one is computed.
// ODD: odd source widths.  Rows then start alternately on and off a 16-byte boundary.  Every staged row gets its own
// column shift d in {0, 1} -- source column c of source row sr sits at staged column c - sc0 + d,
// d = (element offset of the image + sr - sc0) & 1 -- so that the row minus one edge element IS a 16-byte aligned
// bulk copy at both ends; the edge element (first or last of the row) follows as one 8-byte cp.async, and the one
// staged column next to the row that another unit's shift may have used is cleared.  Whole rows only (c_lo = 0,
// c_hi = ws: true for every mode).  The window loop walks even and odd staged rows through two pointers.
template <int HK, int NBUF, int TI, bool ODD>
__global__ void __launch_bounds__(kDirectThreads, 2) direct_bulk_kernel(const __grid_constant__ DirectParams p) {
  unsigned long long* bars = reinterpret_cast<unsigned long long*>(b2c_smem_direct);   // [2]
  double* kflip = reinterpret_cast<double*>(b2c_smem_direct + 16);   // [wk][hk_pad]: column-major, flipped
  constexpr int hk = HK, hk_pad = (HK + 1) & ~1;
  double* S0 = kflip + (size_t)p.wk * hk_pad;
  const int pitch = p.sp_w;
  const size_t src_sz = (size_t)p.hs * p.ws, ker_sz = (size_t)p.hk * p.wk, dst_sz = (size_t)p.hd * p.wd;
  const int group = p.group;
  const long units = group > 1 ? (p.n_images + group - 1) / group : p.n_images * p.tiles_per_image;
  // staged rows per image slot: whole blocks of TI output rows, so that no window has to be clamped (the rows past the
  // tile are zeros and feed only outputs that are not stored)
  const int rows_max = (p.tile_h + TI - 1) / TI * TI + hk - 1, nelem = rows_max * pitch;
  const int bufsz = group * nelem;
  const int sc0 = p.off_w - (p.wk - 1);
  const int c_lo = max(0, sc0), c_hi = min(p.ws, sc0 + p.wd + p.wk - 1);   // copied columns of the source
  const unsigned row_bytes = (unsigned)(c_hi - c_lo) * 8u;
  const int v_lo = c_lo - sc0 + p.col_shift;                                // their first staged column
  for (int i = threadIdx.x; i < NBUF * bufsz; i += kDirectThreads) S0[i] = 0.0;   // halo (and interior, once)
  bulk::fence_async_proxy();   // writers fence, then a barrier, then the copy engine may write
  if (threadIdx.x == 0) {
    bulk::mbar_init(&bars[0], 1);
    bulk::mbar_init(&bars[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  // geometry of a work unit and the copies that stage it into buffer `buf` (all threads call this together)
  auto stage = [&](long work, int buf) {
    const long img0 = group > 1 ? work * group : work / p.tiles_per_image;
    const int ng = group > 1 ? (int)min((long)group, p.n_images - img0) : 1;
    const int i0 = group > 1 ? 0 : (int)(work - img0 * p.tiles_per_image) * p.tile_h;
    const int rows = min(p.tile_h, p.hd - i0) + hk - 1;
    const int sr0 = i0 + p.off_h - (hk - 1);
    const int r_lo = max(0, sr0), r_hi = min(p.hs, sr0 + rows);   // copied rows of the source
    const int nr = max(0, r_hi - r_lo);
    double* S = S0 + buf * bufsz;
    if (!ODD && p.wrap_h > 0) {
      // CIRCULAR_SAME (even widths): every staged row is a source row, (sr0 + u) mod hs, moved whole by one bulk copy;
      // the wk - 1 columns in front of it are the END of the same row and follow as 8-byte cp.asyncs.  Nothing is zero.
      const int L = p.wk - 1;
      if (threadIdx.x == 0) bulk::mbar_expect_tx(&bars[buf], (unsigned)p.ws * 8u * (unsigned)(rows * ng));
      for (int i = threadIdx.x; i < rows * ng; i += kDirectThreads) {
        const int g = i / rows, u = i - g * rows;
        const int r = wrap_index(sr0 + u, p.wrap_h);
        const double* srow = p.src + ((size_t)(img0 + g) * src_sz + (size_t)r * p.ws);
        double* row = S + g * nelem + u * pitch + L + p.col_shift;   // staged address of source column 0
        bulk::bulk_g2s(row, srow, (unsigned)p.ws * 8u, &bars[buf]);
        for (int m = 0; m < L; ++m) cp_async8(row - L + m, srow + p.ws - L + m, 8);
      }
      cp_async_commit();   // one group per unit: the wait below leaves the NEXT unit's group in flight
      return;
    }
    if (p.tiles_per_image > 1 && gridDim.x % p.tiles_per_image != 0) {
      // row tiles: the rows above / below the source change from tile to tile and may hold an earlier tile's data
      // (not when the grid is a multiple of the tiles per image: a CTA then meets the same tile of every image, its
      // zero rows never move, and the fence below -- MEMBAR.ALL.GPU + FENCE.VIEW.ASYNC, 6 % of the warp samples at
      // 100 x 100 (*) 7 x 7 -- is not needed; the host sizes the grid that way)
      const int top = r_lo - sr0, bot = r_hi - sr0;   // staged rows [top, bot) are copied
      for (int i = threadIdx.x; i < (top + rows_max - bot) * pitch; i += kDirectThreads) {
        const int u = i / pitch, v = i - u * pitch;
        S[(u < top ? u : bot + (u - top)) * pitch + v] = 0.0;
      }
      bulk::fence_async_proxy();   // fenced before the barrier that precedes the next copies into this buffer
    }
    // (copies may complete before the byte count is posted: the phase cannot end before thread 0 has arrived)
    if (ODD) {
      if (threadIdx.x == 0) bulk::mbar_expect_tx(&bars[buf], (unsigned)(p.ws - 1) * 8u * (unsigned)(nr * ng));
      for (int i = threadIdx.x; i < nr * ng; i += kDirectThreads) {
        const int g = i / nr, r = r_lo + (i - g * nr);
        const size_t e = (size_t)(img0 + g) * src_sz + (size_t)r * p.ws;   // element offset of the row in the batch
        const int par = (int)(e & 1);                                      // the row starts 8 bytes off a 16-byte boundary
        const int dlt = par ^ (sc0 & 1);                                   // this row's column shift
        double* row = S + g * nelem + (r - sr0) * pitch - sc0 + dlt;       // staged address of source column 0
        const double* srow = p.src + e;
        row[dlt ? -1 : p.ws] = 0.0;                                        // the edge column the other shift would use
        cp_async8(row + (par ? 0 : p.ws - 1), srow + (par ? 0 : p.ws - 1), 8);
        bulk::bulk_g2s(row + par, srow + par, (unsigned)(p.ws - 1) * 8u, &bars[buf]);
      }
      cp_async_commit();
      return;
    }
    if (threadIdx.x == 0) bulk::mbar_expect_tx(&bars[buf], row_bytes * (unsigned)(nr * ng));
    const double* a = p.src + (size_t)img0 * src_sz;
    for (int i = threadIdx.x; i < nr * ng; i += kDirectThreads) {
      const int g = i / nr, r = r_lo + (i - g * nr);
      bulk::bulk_g2s(S + g * nelem + (r - sr0) * pitch + v_lo, a + (size_t)g * src_sz + (size_t)r * p.ws + c_lo, row_bytes, &bars[buf]);
    }
  };

  unsigned phase0 = 0u, phase1 = 0u;
  bool kernel_loaded = false;
  long it = 0;
  if (NBUF == 2 && (long)blockIdx.x < units) stage(blockIdx.x, 0);
  for (long work = blockIdx.x; work < units; work += gridDim.x, ++it) {
    const int buf = NBUF == 2 ? (int)(it & 1) : 0;
    const long img0 = group > 1 ? work * group : work / p.tiles_per_image;
    const int ng = group > 1 ? (int)min((long)group, p.n_images - img0) : 1;
    const int i0 = group > 1 ? 0 : (int)(work - img0 * p.tiles_per_image) * p.tile_h;
    const int th = min(p.tile_h, p.hd - i0);
    __syncthreads();   // the previous unit's readers are done with the other buffer (NBUF 2) / this buffer (NBUF 1) and kflip
    bool staged_next = false;
    if (NBUF == 2) {
      if (work + gridDim.x < units) { stage(work + gridDim.x, buf ^ 1); staged_next = true; }
    } else {
      stage(work, 0);
    }
    const bool load_kernel = !(p.kernel_is_shared && kernel_loaded);
    if (load_kernel) {
      const double* k = p.ker + (p.kernel_is_shared ? 0 : (size_t)img0 * ker_sz);
      for (int idx = threadIdx.x; idx < p.wk * hk_pad; idx += kDirectThreads) {
        const int c = idx / hk_pad, r = idx - c * hk_pad;
        kflip[idx] = r < hk ? __ldg(&k[(size_t)(hk - 1 - r) * p.wk + (p.wk - 1 - c)]) : 0.0;
      }
      kernel_loaded = true;
    }
    bulk::mbar_wait(&bars[buf], buf ? phase1 : phase0);   // (completes at once when the unit has no source rows)
    if (buf) phase1 ^= 1u; else phase0 ^= 1u;
    if (ODD || p.wrap_h > 0) {                                     // this thread's edge / wrapped elements of THIS unit
      if (staged_next) cp_async_wait_group<1>(); else cp_async_wait_group<0>();
    }
    if (ODD || p.wrap_h > 0 || load_kernel) __syncthreads();      // they, cleared columns and kflip were written by other threads
    const double* S = S0 + buf * bufsz;
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
      const double* col = S + (g * nelem + r0 * pitch + j + p.col_shift);
      // ODD: staged row r0 + u carries the column shift (k0 + u) & 1 (see stage): even rows through pe, odd through po
      const int k0 = ODD ? (int)(((size_t)(img0 + g) * src_sz + (size_t)(i0 + p.off_h - (hk - 1) + r0 - (p.off_w - (p.wk - 1)))) & 1) : 0;
      const double* pe = col + k0;
      const double* po = col + pitch + (1 - k0);
      const int pitch2 = 2 * pitch;
#pragma unroll 2
      for (int kc = 0; kc < p.wk; ++kc) {
        double w[TI + HK - 1];
#pragma unroll
        for (int u = 0; u < TI + HK - 1; ++u) w[u] = ODD ? ((u & 1) ? po : pe)[(u >> 1) * pitch2 + kc] : col[u * pitch + kc];
        const double* kv = kflip + kc * hk_pad;
#pragma unroll
        for (int kr = 0; kr < HK; ++kr) {
          const double kk = kv[kr];
#pragma unroll
          for (int t = 0; t < TI; ++t) acc[t] = fma(kk, w[t + kr], acc[t]);
        }
      }
      double* o = d + (size_t)g * dst_sz + (size_t)(i0 + r0) * p.wd + j;
#pragma unroll
      for (int t = 0; t < TI; ++t)
        if (r0 + t < th) o[(size_t)t * p.wd] = acc[t];
    }
  }
}

template <int NBUF, int TI, bool ODD>
static inline direct_fn direct_bulk_pick(int hk) {
  switch (hk) {
    case 1: return direct_bulk_kernel<1, NBUF, TI, ODD>;   case 2: return direct_bulk_kernel<2, NBUF, TI, ODD>;
    case 3: return direct_bulk_kernel<3, NBUF, TI, ODD>;   case 4: return direct_bulk_kernel<4, NBUF, TI, ODD>;
    case 5: return direct_bulk_kernel<5, NBUF, TI, ODD>;   case 6: return direct_bulk_kernel<6, NBUF, TI, ODD>;
    case 7: return direct_bulk_kernel<7, NBUF, TI, ODD>;   case 8: return direct_bulk_kernel<8, NBUF, TI, ODD>;
    case 9: return direct_bulk_kernel<9, NBUF, TI, ODD>;   case 10: return direct_bulk_kernel<10, NBUF, TI, ODD>;
    case 11: return direct_bulk_kernel<11, NBUF, TI, ODD>; case 12: return direct_bulk_kernel<12, NBUF, TI, ODD>;
    case 13: return direct_bulk_kernel<13, NBUF, TI, ODD>; case 14: return direct_bulk_kernel<14, NBUF, TI, ODD>;
    case 15: return direct_bulk_kernel<15, NBUF, TI, ODD>; case 16: return direct_bulk_kernel<16, NBUF, TI, ODD>;
  }
  return nullptr;
}
template <bool ODD>
static inline direct_fn direct_bulk_variant_o(int hk, int nbuf, int ti) {
  switch (ti) {
    case 8: return nbuf == 2 ? direct_bulk_pick<2, 8, ODD>(hk) : direct_bulk_pick<1, 8, ODD>(hk);
    case 10: return nbuf == 2 ? direct_bulk_pick<2, 10, ODD>(hk) : direct_bulk_pick<1, 10, ODD>(hk);
    case 12: return nbuf == 2 ? direct_bulk_pick<2, 12, ODD>(hk) : direct_bulk_pick<1, 12, ODD>(hk);
    case 1: return nbuf == 2 ? direct_bulk_pick<2, 1, ODD>(hk) : direct_bulk_pick<1, 1, ODD>(hk);
  }
  return nullptr;
}
}  // namespace b2c
