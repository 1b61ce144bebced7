// Strip-mode fused FFT convolution with a COMPILE-TIME radix schedule: the tuned form of
// conv_big2d.cu for the period of BASELINE config 3 (100x100 (*) 100x100 CIRCULAR_FULL -> 200 x 200,
// per-image kernels).  Replaces fftw_circular_convolution + crop (reference convolution_fftw.h:169-284).
//
// Same three phases as conv_big2d.cu (rows -> half spectra -> column strips with the product -> rows)
// through a per-CTA slab that stays in L2, with these differences:
//   * 200 = 10 x 20: two passes per transform (prime-factor butterflies of 10 and 20 points in
//     registers) instead of three radix-5/5/8 passes -- the transforms are bound by trips through
//     shared memory and by barriers, not by the FP64 pipe; the spectrum order is a closed formula
//     (no permutation table loads);
//   * the 32 lanes of a warp walk adjacent LINES (rows of a chunk / columns of a strip) and a warp
//     owns one butterfly index: block, offset and twiddles are warp-uniform, the twiddles are
//     constant-bank operands of the FP64 instructions;
//   * column phase: last forward pass of a source column and of its kernel column, the pointwise
//     product and the first backward pass are ONE pass over a block of 20 points in registers;
//   * the slab is strip-major and in place: the block of a strip holds its S rows then its K rows,
//     contiguous, and the strip's result Y replaces it.  A strip moves with bulk asynchronous copies
//     (cp.async.bulk + mbarrier: the copy engine, no registers, no per-thread round trips to L2);
//   * source / kernel reads and result writes carry evict-first hints so that the slabs
//     (148 x 323 KB) stay resident in the 126 MB L2.
#include "bulk_copy.cuh"
#include "launch_cache.h"
#include "conv_params.h"
#include "tw200_table.h"

namespace b2c {

extern __shared__ __align__(16) unsigned char b2c_smem_bigs[];

namespace bigs {

#ifndef B2C_BIGS_TRACE
#define B2C_BIGS_TRACE 0
#endif
#if B2C_BIGS_TRACE
__device__ long long g_acc[16];   // debug variant: cycles accumulated per phase by CTA 0 / thread 0
#define B2C_T(slot) do { if (blockIdx.x == 0 && threadIdx.x == 0) { long long c_; asm volatile("mov.u64 %0, %%clock64;" : "=l"(c_)); g_acc[slot] += c_ - tprev; tprev = c_; } } while (0)
#else
#define B2C_T(slot)
#endif

using namespace bulk;   // bulk_copy.cuh

#ifndef B2C_BIGS_R0
#define B2C_BIGS_R0 20     // first (twiddled) radix of the two-pass 200-point schedule (20 x 10: 1.74 M conv/s, 10 x 20: 1.63 M)
#endif
#ifndef B2C_BIGS_NT
#define B2C_BIGS_NT 192    // threads per CTA in the two-CTAs-per-SM shape
#endif

template <int R> using Rbfly = BigBfly<R>;

template <int P> struct TwTable;
template <> struct TwTable<200> { static __device__ __forceinline__ cplx at(int k) { return c_tw200[k]; } };

// One radix-R stage of a length-P transform on `nlines` lines (element k of line l at W[l*ls + k*es]),
// block length N, lanes = lines, warps = butterfly indices.
template <int P, int R, int N, bool INV>
__device__ __forceinline__ void stage(cplx* W, int es, int ls, int nlines, int warp, int nwarps, int lane,
                                      int nsplit = 1 << 30, int kjump = 0) {
  constexpr int M = N / R, PER_LINE = P / R, TSTEP = P / N;
  // 32 lines per warp task, or -- for strips of at most 16 lines -- two butterflies of 16 lines each
  const int lpt = nlines <= 8 ? 8 : nlines <= 16 ? 16 : 32, subs = 32 / lpt;
  const int ntrips = (nlines + lpt - 1) / lpt, total = PER_LINE * ntrips;
  for (int wt = warp; wt * subs < total; wt += nwarps) {
    const int task = wt * subs + lane / lpt;
    const int t = task / ntrips, line = (task - t * ntrips) * lpt + lane % lpt;
    const int b = t / M, j = t - b * M;
    cplx* base = W + (b * N + j) * es;
    if (task < total && line < nlines) {
      cplx* p = base + line * ls + (line >= nsplit ? kjump : 0);   // lines past nsplit live in a second array
      cplx v[R];
#pragma unroll
      for (int q = 0; q < R; ++q) v[q] = p[q * M * es];
      if (INV) {
        if (M > 1) {
#pragma unroll
          for (int q = 1; q < R; ++q) v[q] = cmulc(v[q], TwTable<P>::at(j * q * TSTEP));
        }
#pragma unroll
        for (int q = 0; q < R; ++q) v[q] = cswap(v[q]);
        Rbfly<R>::run(v);
#pragma unroll
        for (int q = 0; q < R; ++q) v[q] = cswap(v[q]);
      } else {
        Rbfly<R>::run(v);
        if (M > 1) {
#pragma unroll
          for (int q = 1; q < R; ++q) v[q] = cmul(v[q], TwTable<P>::at(j * q * TSTEP));
        }
      }
#pragma unroll
      for (int q = 0; q < R; ++q) p[q * M * es] = v[q];
    }
  }
  __syncthreads();
}

// Two-pass schedule P = R0 * R1: a radix-R0 pass with twiddles, then a twiddle-free radix-R1 pass (each a
// prime-factor butterfly in registers).  Two passes instead of three radix-5/5/8 passes: the strip
// transforms are bound by shared-memory traffic, not by the FP64 pipe.  Frequency f ends at position
// (f % R0) * R1 + f / R0.
template <int P, int R0, int R1>
struct Fft {
  static_assert(R0 * R1 == P, "two-pass schedule");
  static __device__ __forceinline__ int pos_of(int f) { return (f % R0) * R1 + f / R0; }
  static __device__ __forceinline__ void forward(cplx* W, int es, int ls, int nlines, int warp, int nwarps, int lane) {
    stage<P, R0, P, false>(W, es, ls, nlines, warp, nwarps, lane);
    stage<P, R1, R1, false>(W, es, ls, nlines, warp, nwarps, lane);
  }
  static __device__ __forceinline__ void backward(cplx* W, int es, int ls, int nlines, int warp, int nwarps, int lane) {
    stage<P, R1, R1, true>(W, es, ls, nlines, warp, nwarps, lane);
    stage<P, R0, P, true>(W, es, ls, nlines, warp, nwarps, lane);
  }
  // Column phase, fused: (last forward pass of line l and of its partner koff elements further) ->
  // pointwise product * scale -> first backward pass, all on one block of R1 points in registers
  // (convolution_fftw.h:195-204, 210-211 between the two transforms).  One barrier and two trips
  // through shared memory fewer than forward-pass / product / backward-pass.
  static __device__ __forceinline__ void middle(cplx* W, int es, int ls, int nlines, int warp, int nwarps, int lane,
                                                int koff, double scale) {
    const int lpt = nlines <= 8 ? 8 : nlines <= 16 ? 16 : 32, subs = 32 / lpt;
    const int ntrips = (nlines + lpt - 1) / lpt, total = R0 * ntrips;
    for (int wt = warp; wt * subs < total; wt += nwarps) {
      const int task = wt * subs + lane / lpt;
      const int b = task / ntrips, line = (task - b * ntrips) * lpt + lane % lpt;
      if (task < total && line < nlines) {
        cplx* ps = W + b * R1 * es + line * ls;
        cplx k[R1], v[R1];
#pragma unroll
        for (int q = 0; q < R1; ++q) k[q] = ps[q * es + koff];
        Rbfly<R1>::run(k);
#pragma unroll
        for (int q = 0; q < R1; ++q) v[q] = ps[q * es];
        Rbfly<R1>::run(v);
#pragma unroll
        for (int q = 0; q < R1; ++q) {
          const cplx y = cmul(v[q], k[q]);
          v[q] = make_double2(y.y * scale, y.x * scale);   // swapped: inverse butterfly = forward on (im, re)
        }
        Rbfly<R1>::run(v);
#pragma unroll
        for (int q = 0; q < R1; ++q) ps[q * es] = cswap(v[q]);
      }
    }
    __syncthreads();
  }
  static __device__ __forceinline__ void forward_first(cplx* W, int es, int ls, int nlines, int warp, int nwarps, int lane,
                                                       int nsplit, int kjump) {
    stage<P, R0, P, false>(W, es, ls, nlines, warp, nwarps, lane, nsplit, kjump);
  }
  static __device__ __forceinline__ void backward_last(cplx* W, int es, int ls, int nlines, int warp, int nwarps, int lane) {
    stage<P, R0, P, true>(W, es, ls, nlines, warp, nwarps, lane);
  }
};

__device__ __forceinline__ double wrapped_tap(const BigParams& p, const double* ker, int r, int c) {
  double acc = 0.0;   // taps congruent modulo the period add up (convolution_fftw.h:185-187)
  for (int i = r; i < p.hk; i += p.kwrap_h)
    for (int j = c; j < p.wk; j += p.kwrap_w) acc += ker[(size_t)i * p.wk + j];
  return acc;
}

template <int P, int R0, int R1, int NT, int CTAS, int CS>
__global__ void __launch_bounds__(NT, CTAS) conv_kernel(const __grid_constant__ BigParams p) {
  typedef Fft<P, R0, R1> F;
  cplx* W = reinterpret_cast<cplx*>(b2c_smem_bigs);
  const int lane = threadIdx.x & 31;
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);   // uniform for the compiler
  const int nwarps = blockDim.x >> 5;
  const int hke = min(p.hk, p.kwrap_h), wke = min(p.wk, p.kwrap_w);
  const int rows_in = max(p.hs, hke);
  const int HW = p.half_w, ldw = p.ldw;
  const bool plain_kernel = (p.hk <= p.kwrap_h) && (p.wk <= p.kwrap_w);
  const size_t src_sz = (size_t)p.src_h * p.src_w, dst_sz = (size_t)p.hd * p.wd;
  const bool shifted = (p.sh_h | p.sh_w) != 0;   // periodically extended source (CIRCULAR_SAME by linear convolution)
  // Per-CTA slab, strip-major: the block of strip v0 starts at rows_total * v0 and holds [hs][nc] S rows
  // followed by [hke][nc] K rows -- one contiguous run per transfer; the strip's Y[rows_y][nc] later
  // replaces the block in place.
  constexpr int NQ = 1;                                      // pieces per bulk transfer (more pieces measured no faster)
  const int rows_total = max(p.hs + hke, p.rows_y);
  cplx* slab = p.slab + (size_t)blockIdx.x * (size_t)rows_total * HW;
  __shared__ unsigned long long bar_storage;
  __shared__ double row_unscale[64];   // per row of a chunk: 0.5 / (power-of-two scale of its kernel row)
  unsigned long long* bar = &bar_storage;
  unsigned bar_phase = 0;
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  fence_async_proxy();
  __syncthreads();

#if B2C_BIGS_TRACE
  long long tprev; asm volatile("mov.u64 %0, %%clock64;" : "=l"(tprev));
#endif
  for (long img = blockIdx.x; img < p.n_images; img += gridDim.x) {
    const double* a = p.src + (size_t)img * src_sz;
    const double* k = p.ker + (size_t)img * p.ker_stride;

    // ---------------- phase R: z_r = s_r + i*k_r, forward rows, Hermitian split -> slab
    for (int r0 = 0; r0 < rows_in; r0 += p.rows_chunk) {
      const int nr = min(p.rows_chunk, rows_in - r0);
      if (p.bulk_rows) {
        // The copy engine brings the chunk's source rows and kernel rows into the row buffer itself -- source row at the
        // first P doubles of its buffer row, kernel row at the next P -- all rows in flight at once; each warp then
        // expands its rows in place (reads into registers, __syncwarp, writes).  With one row per warp loaded by
        // ordinary loads the phase sat on one HBM round trip per row: 45 k of 320 k cycles per image.
        const int ns = max(0, min(nr, p.hs - r0)), nk = max(0, min(nr, hke - r0));
        if (threadIdx.x == 0) mbar_expect_tx(bar, (unsigned)((ns * p.ws + nk * p.wk) * sizeof(double)));
        for (int i = threadIdx.x; i < ns + nk; i += NT) {
          const bool isk = i >= ns;
          const int rr = isk ? i - ns : i;
          double* to = reinterpret_cast<double*>(W + rr * ldw) + (isk ? P : 0);
          if (isk) bulk_g2s(to, k + (size_t)(r0 + rr) * p.wk, (unsigned)(p.wk * sizeof(double)), bar);
          else bulk_g2s(to, a + (size_t)(r0 + rr) * p.ws, (unsigned)(p.ws * sizeof(double)), bar);
        }
        mbar_wait(bar, bar_phase);
        bar_phase ^= 1u;
        B2C_T(14);
      }
      for (int rr = warp; rr < nr; rr += nwarps) {
        const int r = r0 + rr;
        cplx* wrow = W + rr * ldw;
        const double* arow = a + (size_t)r * p.ws;
        const double* krow = k + (size_t)r * p.wk;
        constexpr int NCH = (P + 31) / 32;       // all loads of a row are in flight before the first store
        cplx v[NCH];
        if (p.bulk_rows) {
          const double* srow = reinterpret_cast<const double*>(wrow);
#pragma unroll
          for (int u = 0; u < NCH; ++u) {
            const int c = lane + 32 * u;
            v[u] = make_double2(0.0, 0.0);
            if (r < p.hs && c < p.ws) v[u].x = srow[c];
            if (r < hke && c < wke) v[u].y = srow[P + c];
          }
          __syncwarp();   // every lane has read the staged row before any lane overwrites it
        } else {
#pragma unroll
        for (int u = 0; u < NCH; ++u) {
          const int c = lane + 32 * u;
          v[u] = make_double2(0.0, 0.0);
          if (r < p.hs && c < p.ws) {
            if (shifted) {
              int sr = r - p.sh_h, sc = c - p.sh_w;   // shifts are below the source extents: one wrap
              if (sr < 0) sr += p.src_h; else if (sr >= p.src_h) sr -= p.src_h;
              if (sc < 0) sc += p.src_w; else if (sc >= p.src_w) sc -= p.src_w;
              v[u].x = __ldg(a + (size_t)sr * p.src_w + sc);
            } else {
              v[u].x = __ldcs(arow + c);          // streamed once: evict first, keep the slabs in L2
            }
          }
          if (r < hke && c < wke) v[u].y = plain_kernel ? __ldcs(krow + c) : wrapped_tap(p, k, r, c);
        }
        }
        // Source row and kernel row share one transform and are separated again right after it: the kernel
        // row is scaled to the source row's magnitude by a power of two (fft_device.cuh: balance_from_maxima;
        // the reference transforms them separately, convolution_fftw.h:190-191), undone at the split.
        double ms = 0.0, mk = 0.0;
#pragma unroll
        for (int u = 0; u < NCH; ++u) { ms = fmax(ms, fabs(v[u].x)); mk = fmax(mk, fabs(v[u].y)); }
        // row maxima by their exponent words (one warp-reduce instruction each); alpha = 2^(exponent difference)
        const unsigned es = __reduce_max_sync(0xffffffffu, (unsigned)__double2hiint(ms)) >> 20;
        const unsigned ek = __reduce_max_sync(0xffffffffu, (unsigned)__double2hiint(mk)) >> 20;
        const bool usable = es > 0 && ek > 0 && es < 2047 && ek < 2047;   // zero / subnormal / inf / nan rows: no scaling
        const double al = usable ? __hiloint2double((int)((1023u + es - ek) << 20), 0) : 1.0;
        if (lane == 0) row_unscale[rr] = 0.5 / al;
#pragma unroll
        for (int u = 0; u < NCH; ++u)
          if (lane + 32 * u < P) wrow[lane + 32 * u] = make_double2(v[u].x, v[u].y * al);
      }
      __syncthreads();
      B2C_T(0);
      F::forward(W, 1, ldw, nr, warp, nwarps, lane);
      B2C_T(1);
      // S(v) = (Z(v) + conj Z(-v)) / 2,  K(v) = -i (Z(v) - conj Z(-v)) / 2
      for (int rr = warp; rr < nr; rr += nwarps) {
        const int r = r0 + rr;
        const cplx* wrow = W + rr * ldw;
        const double hk2 = row_unscale[rr];     // 0.5 / alpha of this row
        for (int v = lane; v < HW; v += 32) {
          const cplx z1 = wrow[F::pos_of(v)];
          const cplx z2 = wrow[F::pos_of(v ? P - v : 0)];
          const int v0 = v & ~(CS - 1), nc = min(CS, HW - v0);
          cplx* blk = slab + (size_t)rows_total * v0 + (v - v0);
          if (r < p.hs) blk[r * nc] = make_double2(0.5 * (z1.x + z2.x), 0.5 * (z1.y - z2.y));
          if (r < hke) blk[(p.hs + r) * nc] = make_double2(hk2 * (z1.y + z2.y), hk2 * (z2.x - z1.x));
        }
      }
      __threadfence();          // slab rows must be in L2 and ordered before the copy engine reads them
      fence_async_proxy_global();
      fence_async_proxy();      // and the row buffer is next written by the copy engine
      __syncthreads();
      B2C_T(2);
    }

    // ---------------- phase C: strips of half-spectrum columns: both column transforms, product, backward
    for (int v0 = 0; v0 < HW; v0 += CS) {
      const int nc = min(CS, HW - v0);
      cplx* WK = W + P * nc;                                  // S columns in W[r][nc], K columns in WK[r][nc]
      const cplx* blk = slab + (size_t)rows_total * v0;       // this strip's block: [hs][nc] S rows, then [hke][nc] K rows
      // Two bulk asynchronous copies (the copy engine, no registers, no per-thread round trips to L2)
      // while the threads clear the zero-padding rows.
      if (threadIdx.x == 0) mbar_expect_tx(bar, (unsigned)((p.hs + hke) * nc * sizeof(cplx)));
      if (threadIdx.x < 2 * NQ) {
        const int q = threadIdx.x % NQ, isk = threadIdx.x / NQ;
        const int rows = isk ? hke : p.hs, per = (rows + NQ - 1) / NQ, ra = min(rows, q * per), rb = min(rows, ra + per);
        if (rb > ra)
          bulk_g2s((isk ? WK : W) + ra * nc, blk + (size_t)((isk ? p.hs : 0) + ra) * nc, (unsigned)((rb - ra) * nc * sizeof(cplx)), bar);
      }
      for (int i = p.hs * nc + threadIdx.x; i < P * nc; i += NT) W[i] = make_double2(0.0, 0.0);
      for (int i = hke * nc + threadIdx.x; i < P * nc; i += NT) WK[i] = make_double2(0.0, 0.0);
      mbar_wait(bar, bar_phase);
      bar_phase ^= 1u;
      __syncthreads();
      B2C_T(3);
      const int kjump = P * nc - nc;                          // line nc + c is column c of WK
      F::forward_first(W, nc, 1, 2 * nc, warp, nwarps, lane, nc, kjump);
      B2C_T(4);
      F::middle(W, nc, 1, nc, warp, nwarps, lane, P * nc, p.scale);
      B2C_T(5);
      F::backward_last(W, nc, 1, nc, warp, nwarps, lane);
      B2C_T(6);
      fence_async_proxy();      // the strip was written by ordinary stores; the copy engine reads it next
      __syncthreads();
      B2C_T(11);
      if (threadIdx.x < NQ) {   // Y[rows_y][nc] replaces the strip's S|K block
        const int per = (p.rows_y + NQ - 1) / NQ, ra = min(p.rows_y, (int)threadIdx.x * per), rb = min(p.rows_y, ra + per);
        if (rb > ra)
          bulk_s2g(slab + (size_t)rows_total * v0 + (size_t)ra * nc, W + ra * nc, (unsigned)((rb - ra) * nc * sizeof(cplx)));
        bulk_commit();
        bulk_wait_read();       // the buffer may be reused; the write to the slab completes in the background
      }
      B2C_T(12);
      __syncthreads();
      B2C_T(7);
    }
    if (threadIdx.x < NQ) bulk_wait_all();   // every Y block has landed before the row phase reads it
    __syncthreads();
    B2C_T(13);

    // ---------------- phase R': W = Y_r + i*Y_{r+1} (Hermitian-extended), backward rows, crop -> dst
    double* d = p.dst + (size_t)img * dst_sz;
    const int pairs = (p.hd + 1) / 2;
    for (int q0 = 0; q0 < pairs; q0 += p.rows_chunk) {
      const int nq = min(p.rows_chunk, pairs - q0);
      if (p.bulk_rows) {
        // The chunk's rows of Y travel by bulk copies as well: the slab is strip-major, so a row is one piece of
        // nc * 16 bytes per strip; rows 2q and 2q+1 land side by side in buffer row q (HW complex each) and are
        // combined in place below.
        constexpr int NS = (P / 2 + 1 + CS - 1) / CS;    // strips
        const int nrows = min(2 * nq, p.hd - 2 * q0);
        if (threadIdx.x == 0) mbar_expect_tx(bar, (unsigned)(nrows * HW * sizeof(cplx)));
        for (int i = threadIdx.x; i < nrows * NS; i += NT) {
          const int ri = i / NS, s = i - ri * NS, v0 = s * CS, nc = min(CS, HW - v0);
          const int row = p.off_h + 2 * q0 + ri;
          bulk_g2s(W + (ri >> 1) * ldw + (ri & 1) * HW + v0, slab + (size_t)rows_total * v0 + (size_t)row * nc,
                   (unsigned)(nc * sizeof(cplx)), bar);
        }
        mbar_wait(bar, bar_phase);
        bar_phase ^= 1u;
        B2C_T(15);
      }
      for (int qq = warp; qq < nq; qq += nwarps) {
        const int i0 = 2 * (q0 + qq), i1 = i0 + 1;
        const int row0 = p.off_h + i0, row1 = p.off_h + i1;
        const bool has1 = i1 < p.hd;
        cplx* wrow = W + qq * ldw;
        constexpr int NCH = (P + 31) / 32;
        cplx y0[NCH], y1[NCH];
        int pos[NCH];
#pragma unroll
        for (int u = 0; u < NCH; ++u) {
          const int v = lane + 32 * u;
          pos[u] = -1;
          if (v < P) {
            const int vv = v < HW ? v : P - v;
            if (p.bulk_rows) {
              y0[u] = wrow[vv];
              y1[u] = has1 ? wrow[HW + vv] : make_double2(0.0, 0.0);
            } else {
              const int v0 = vv & ~(CS - 1), nc = min(CS, HW - v0);
              const cplx* blk = slab + (size_t)rows_total * v0 + (vv - v0);
              y0[u] = __ldcg(blk + row0 * nc);   // written by the copy engine: read at L2, never from a stale L1 line
              y1[u] = has1 ? __ldcg(blk + row1 * nc) : make_double2(0.0, 0.0);
            }
            if (v >= HW) { y0[u].y = -y0[u].y; y1[u].y = -y1[u].y; }   // Hermitian mirror
            pos[u] = F::pos_of(v);
          }
        }
        if (p.bulk_rows) __syncwarp();   // staged rows are read before the in-place combination overwrites them
#pragma unroll
        for (int u = 0; u < NCH; ++u)
          if (pos[u] >= 0) wrow[pos[u]] = make_double2(y0[u].x - y1[u].y, y0[u].y + y1[u].x);
      }
      __syncthreads();
      B2C_T(8);
      F::backward(W, 1, ldw, nq, warp, nwarps, lane);
      B2C_T(9);
      for (int ri = warp; ri < 2 * nq; ri += nwarps) {
        const int i = 2 * q0 + ri;
        if (i < p.hd) {
          const cplx* wrow = W + (ri >> 1) * ldw + p.off_w;
          double* drow = d + (size_t)i * p.wd;
          if (ri & 1) { for (int j = lane; j < p.wd; j += 32) __stcs(drow + j, wrow[j].y); }
          else        { for (int j = lane; j < p.wd; j += 32) __stcs(drow + j, wrow[j].x); }
        }
      }
      fence_async_proxy();      // the row buffer may next be written by the copy engine
      __syncthreads();
      B2C_T(10);
    }
  }
}

}  // namespace bigs

// true when a compile-time strip kernel exists for this (square) period and radix schedule
#if B2C_BIGS_TRACE
extern "C" int b2c_debug_bigs_read(long long* host) { return (int)cudaMemcpyFromSymbol(host, bigs::g_acc, sizeof(long long) * 16); }
#endif

bool big2ds_has(int Ph, int Pw, const int* rad_h, int ns_h, const int* rad_w, int ns_w) {
  (void)rad_h; (void)ns_h; (void)rad_w; (void)ns_w;   // the kernel carries its own radix schedule and spectrum order
  return Ph == 200 && Pw == 200;
}

// Two shapes: one CTA of 384 threads per SM with 32-column strips and 64-row chunks, or two CTAs of 192
// threads with 16-column strips and 32-row chunks (one copies while the other transforms).  168 registers
// per thread either way: the fused middle pass keeps 2 x 20 complex values.
int big2ds_ctas_per_sm() {
  // measured: one CTA 1.53, two 1.63 M conv/s; three (160 threads) / four (128 threads) CTAs with 8-column strips and
  // 16-row chunks 1.68 / 1.69 against 1.79 for two once the rows were bulk-staged (shapes removed again)
  static const int v = getenv("B2C_BIG_CTAS") ? atoi(getenv("B2C_BIG_CTAS")) : 2;
  return v == 1 ? 1 : 2;
}
// strip width and rows per chunk of each shape
void big2ds_shape(int ctas, int* cols_strip, int* rows_chunk) {
  *cols_strip = ctas == 1 ? 32 : 16;
  *rows_chunk = ctas == 1 ? 64 : 32;
}

template <int NT, int CTAS, int CS>
static cudaError_t launch_shape(const BigParams& p, int grid, size_t smem, cudaStream_t st) {
  auto kern = bigs::conv_kernel<200, B2C_BIGS_R0, 200 / B2C_BIGS_R0, NT, CTAS, CS>;
  ensure_dyn_smem(reinterpret_cast<const void*>(kern), smem);
  kern<<<grid, NT, smem, st>>>(p);
  return cudaGetLastError();
}

cudaError_t launch_conv_big2ds(const BigParams& p, int grid, size_t smem, cudaStream_t st) {
  if (big2ds_ctas_per_sm() == 1) return launch_shape<384, 1, 32>(p, grid, smem, st);
  return launch_shape<B2C_BIGS_NT, 2, 16>(p, grid, smem, st);
}

}  // namespace b2c
