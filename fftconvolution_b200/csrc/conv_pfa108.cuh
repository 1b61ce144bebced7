// The five passes of the period-108 convolution kernel (conv_pfa108.cu), written per thread index so that
// tools/emulate_pfa108.cu can step them on the CPU during development.  Product code reaches them only through
// pfa::conv_kernel.
//
// 108 = 4 x 27 with gcd(4, 27) = 1: on both axes the 108-point DFT is a Good-Thomas (prime-factor) 4 x 27
// two-dimensional DFT with NO twiddle factors between the stages:
//     input  index n = (27*n4 + 4*n27) mod 108      n4 in 0..3, n27 in 0..26
//     output index k = (81*k4 + 28*k27) mod 108     X[k] = sum_{n4,n27} x[n] W4^(n4 k4) W27^(n27 k27)
// so the whole 108 x 108 real transform, the product with the kernel spectrum and the way back take five
// in-place passes over ONE shared buffer Z[54][LD] (image columns c and c + 54 ride one complex line):
//   1 vertical 27-point DFTs straight from global memory          (thread = line, n4)
//   2 vertical 4-point DFTs, Hermitian split of the column pair, horizontal 4-point DFTs
//   3 horizontal 27-point DFT, kernel spectrum, inverse 27-point DFT
//   4 inverse of pass 2
//   5 inverse vertical 27-point DFTs, crop, straight to global memory
// against the seven passes and the per-thread twiddle registers of the (12, 9) Cooley-Tukey member of
// conv_tuned2.cu.  Replaces fftw_circular_convolution + crop (reference convolution_fftw.h:169-284).
//
// Layout.  A line has two regions of 56 slots.  Vertical (group g = k27, 4-point index s) sits at 14 s + g for
// g <= 13 and at 56 + 14 s + (27 - g) for g >= 14, so group u in 1..13 and its mirror 27 - u share the index u
// (vertical frequency kv = (81 s + 28 g) mod 108 has its mirror -kv at (27 - g, (4 - s) mod 4)); group 0 is its own
// mirror and region-B index 0 is spare.  The half spectrum kept is kk = 14 s + u, u in 0..13 (56 entries;
// (u = 0, s = 3) duplicates the mirror of (0, 1)).  Lines: horizontal index n27' owns lines L0 = 4 n27' mod 54
// (columns 4 n27' and + 54) and L1 = (4 n27' + 27) mod 54 (columns 4 n27' + 27 and + 81); pass 2 leaves horizontal
// 4-point output k4' of half-spectrum frequency kk in line L_(k4' & 1), region k4' >> 1, slot kk.
// Every shared access of a quarter-warp is 8 consecutive 16-byte slots or 8 lines at the odd pitch LD: no bank
// conflicts by construction.
#pragma once
#include "fft_device.cuh"

namespace b2c {
namespace pfa {

struct Params {
  int hs, ws, hd, wd, off_h, off_w;
  // periodic-extension form of CIRCULAR_SAME (planner.h): the transform input has hs_in x ws_in positions,
  // position (r, c) holds source sample ((r - sh_h) mod hs, (c - sh_w) mod ws); hs, ws, 0, 0 otherwise
  int hs_in, ws_in, sh_h, sh_w;
  const double* src;
  double* dst;
  long n_images;
  const cplx* kspec;   // [4 k4'][27 k27'][56 kk], pre-scaled by 0.5 / 108^2
};

constexpr int P = 108, HALF = 54, LD = 113, NU = 14, KK = 56, REGB = 56, THREADS = 224, LPAD = 56;
constexpr size_t SMEM_BYTES = sizeof(cplx) * (size_t)HALF * LD;   // 97 632: two CTAs per SM inside the 196 KB carve-out

// Constants of the 27-point DFT: sin(2 pi / 3), then exp(-2 pi i m / 27) for m = 1..8, 10, 12, 14, 16.  On the device
// they are constant-bank operands of the FP64 instructions (literals would cost two uniform moves each and the
// kernel must stay small: beyond ~40 KB of code the warps of two CTAs starve on instruction fetch).
#define B2C_PFA_CONSTANTS                                                                                          \
  0.8660254037844386467637, 0.9730448705798238388329, -0.2306158707424401784502, 0.8936326403234122481926,         \
  -0.448799180200462172785, 0.7660444431189780352024, -0.6427876096865393263226, 0.5971585917027861648519,         \
  -0.8021231927550437850833, 0.396079766039156823696, -0.918216106880274014759, 0.1736481776669303488517,          \
  -0.9848077530122080593667, -0.05814482891047582853875, -0.9983081582712682080478, -0.2868032327110902531033,     \
  -0.9579895123154888744374, -0.6862416378687335857296, -0.7273736415730486959872, -0.9396926207859083840541,      \
  -0.3420201433256687330441, -0.9932383577419429885479, 0.1160929141252302296757, -0.8354878114129364196538,       \
  0.5495089780708060352628
__constant__ double c_pfa[25] = {B2C_PFA_CONSTANTS};
static const double h_pfa[25] = {B2C_PFA_CONSTANTS};   // host copy, for the CPU emulation of the passes only
#ifdef __CUDA_ARCH__
#define B2C_PFA_K(i) c_pfa[i]
#define B2C_PFA_LDG(ptr) __ldg(ptr)
#else
#define B2C_PFA_K(i) h_pfa[i]
#define B2C_PFA_LDG(ptr) (*(ptr))
#endif
// Predicated global accesses: one instruction each, no branch and no select (a conditional expression compiles to a
// branch per element here, and the passes must stay small and cheap to issue).
B2C_MATH double ldg_if(const double* ptr, bool on) {
#ifdef __CUDA_ARCH__
  double v;
  asm volatile("{\n .reg .pred p;\n setp.ne.u32 p, %2, 0;\n mov.f64 %0, 0d0000000000000000;\n @p ld.global.nc.f64 %0, [%1];\n}"
               : "=d"(v) : "l"(ptr), "r"((unsigned)on));
  return v;
#else
  return on ? *ptr : 0.0;
#endif
}
B2C_MATH void stg_if(double* ptr, double v, bool on) {
#ifdef __CUDA_ARCH__
  asm volatile("{\n .reg .pred p;\n setp.ne.u32 p, %2, 0;\n @p st.global.f64 [%0], %1;\n}" ::"l"(ptr), "d"(v), "r"((unsigned)on) : "memory");
#else
  if (on) *ptr = v;
#endif
}

// table slot of exp(-2 pi i m / 27): m = 1..8 -> 1, 3, .., 15; 10, 12, 14, 16 -> 17, 19, 21, 23
__host__ __device__ constexpr int w27_slot(int m) { return m <= 8 ? 2 * m - 1 : m + 7; }
#define B2C_PFA_W27(m) make_double2(B2C_PFA_K(w27_slot(m)), B2C_PFA_K(w27_slot(m) + 1))

B2C_MATH void dft3(cplx& a, cplx& b, cplx& c) {
  const double s = B2C_PFA_K(0);
  const cplx t = cadd(b, c), d = csub(b, c);
  const cplx m = make_double2(fma(-0.5, t.x, a.x), fma(-0.5, t.y, a.y));
  a = cadd(a, t);
  b = make_double2(fma(s, d.y, m.x), fma(-s, d.x, m.y));
  c = make_double2(fma(-s, d.y, m.x), fma(s, d.x, m.y));
}

// Forward 27-point DFT in registers, natural order in and out: 27 = 3 x 3 x 3 Cooley-Tukey with the constants
// above; n = 9 a + b, k = ka + 3 kb: DFT3 over a, twiddle W27^(b ka), then a 9-point DFT over b (itself 3 x 3 with
// W9 = W27^3).  436 FP64 instructions.
B2C_MATH void dft27(cplx* v) {
#pragma unroll
  for (int b = 0; b < 9; ++b) {
    dft3(v[b], v[b + 9], v[b + 18]);
    if (b) {
      v[b + 9] = cmul(v[b + 9], B2C_PFA_W27(b));
      v[b + 18] = cmul(v[b + 18], B2C_PFA_W27(2 * b));
    }
  }
  cplx o[27];
#pragma unroll
  for (int ka = 0; ka < 3; ++ka) {
    cplx* x = v + 9 * ka;   // 9-point DFT of x[0..8] -> o[ka + 3 kb]
#pragma unroll
    for (int c = 0; c < 3; ++c) dft3(x[c], x[c + 3], x[c + 6]);
    x[4] = cmul(x[4], B2C_PFA_W27(3));  x[5] = cmul(x[5], B2C_PFA_W27(6));
    x[7] = cmul(x[7], B2C_PFA_W27(6));  x[8] = cmul(x[8], B2C_PFA_W27(12));
#pragma unroll
    for (int k1 = 0; k1 < 3; ++k1) {
      dft3(x[3 * k1], x[3 * k1 + 1], x[3 * k1 + 2]);
#pragma unroll
      for (int k2 = 0; k2 < 3; ++k2) o[ka + 3 * (k1 + 3 * k2)] = x[3 * k1 + k2];
    }
  }
#pragma unroll
  for (int k = 0; k < 27; ++k) v[k] = o[k];
}

// position of vertical (group g = k27, 4-point index s) inside a line
__host__ __device__ constexpr int vpos(int g, int s) { return g < NU ? NU * s + g : REGB + NU * s + (27 - g); }

// ---- pass 1: vertical 27-point DFTs, global -> Z.  Thread (n4 = j = t / 56, line = t % 56 < 54): element n27 of
// its sequence is row (27 j + 4 n27) mod 108, columns `line` (real part) and `line + 54` (imaginary part).
struct VThread {   // what a thread of passes 1 and 5 keeps across images
  int j, line;
  bool act;
};
// The row numbers of a thread depend only on its n4: left alone, the compiler hoists all 27 of them (and the offsets
// and predicates built on them) out of the image loop and pins ~60 registers for the whole kernel, which starves
// every other pass.  Re-deriving them per image costs a few integer instructions; this hides the invariance.
B2C_MATH int opaque(int x) {
#ifdef __CUDA_ARCH__
  asm volatile("" : "+r"(x));
#endif
  return x;
}
B2C_MATH VThread vthread(int t) {
  VThread q;
  q.j = t / LPAD; q.line = t - q.j * LPAD;
  q.act = q.line < HALF;
  return q;
}

// EXT: the periodic-extension form of CIRCULAR_SAME (its own kernel instantiation: the common form stays small)
template <bool EXT>
B2C_MATH void p1_load(cplx* v, const VThread& q, const Params& p, const double* __restrict__ x) {
  const bool c0 = q.line < p.ws_in, c1 = q.line + HALF < p.ws_in;
  const int j27 = opaque(27 * q.j);
  if (!EXT) {
    const int h0 = c0 ? p.hs : 0, h1 = c1 ? p.hs : 0;   // column tests folded into the row limits
    const unsigned ws = (unsigned)p.ws;
    const double* xc = x + q.line;
#pragma unroll
    for (int n = 0; n < 27; ++n) {
      int r = j27 + 4 * n;
      if (r >= P) r -= P;
      const double* xr = xc + (unsigned)r * ws;
      v[n].x = ldg_if(xr, r < h0);
      v[n].y = ldg_if(xr + HALF, r < h1);
    }
  } else {   // source extended periodically by (sh_h, sh_w) samples in front
    int sc0 = q.line - p.sh_w, sc1 = q.line + HALF - p.sh_w;
    if (sc0 < 0) sc0 += p.ws;
    if (sc1 < 0) sc1 += p.ws;
#pragma unroll
    for (int n = 0; n < 27; ++n) {
      int r = j27 + 4 * n;
      if (r >= P) r -= P;
      int sr = r - p.sh_h;
      if (sr < 0) sr += p.hs;
      const bool in = r < p.hs_in;
      v[n].x = (in && c0) ? B2C_PFA_LDG(x + (unsigned)(sr * p.ws + sc0)) : 0.0;
      v[n].y = (in && c1) ? B2C_PFA_LDG(x + (unsigned)(sr * p.ws + sc1)) : 0.0;
    }
  }
}
B2C_MATH void p1_store(const cplx* v, const VThread& q, cplx* Z) {
  cplx* z = Z + q.line * LD + NU * q.j;
#pragma unroll
  for (int k = 0; k < 27; ++k) z[vpos(k, 0)] = v[k];
}

// ---- pass 5: inverse vertical DFT27, crop, Z -> global.
B2C_MATH void p5_load(cplx* v, const VThread& q, const cplx* Z) {
  const cplx* z = Z + q.line * LD + NU * q.j;
#pragma unroll
  for (int k = 0; k < 27; ++k) v[k] = cswap(z[vpos(k, 0)]);
}
B2C_MATH void p5_store(const cplx* v, const VThread& q, const Params& p, double* __restrict__ y) {
  const int j0 = q.line - p.off_w, j1 = q.line + HALF - p.off_w;
  // output row i = r - off_h must lie in [0, hd); the column tests are folded into the two limits
  const unsigned hd0 = (j0 >= 0 && j0 < p.wd) ? (unsigned)p.hd : 0u, hd1 = (j1 >= 0 && j1 < p.wd) ? (unsigned)p.hd : 0u;
  const unsigned wd = (unsigned)p.wd;
  const int j27 = opaque(27 * q.j);
  double* yc = y + j0;   // only dereferenced where the column test passed
#pragma unroll
  for (int n = 0; n < 27; ++n) {
    int r = j27 + 4 * n;
    if (r >= P) r -= P;
    const unsigned i = (unsigned)(r - p.off_h);   // wraps for rows above the window: fails both tests
    double* yr = yc + i * wd;
    // v is still re/im swapped: .y is column `line`, .x is column `line + 54`
    stg_if(yr, v[n].y, i < hd0);
    stg_if(yr + HALF, v[n].x, i < hd1);
  }
}

struct Pair {   // the two lines and the part flags of horizontal index n27'
  int l0, l1;
  bool f0, f1;   // column 4 n27' (resp. 4 n27' + 27) is the IMAGINARY part of its line
};
B2C_MATH Pair pair_of(int n) {
  Pair q;
  const int c0 = 4 * n;
  int c1 = c0 + 27;
  if (c1 >= P) c1 -= P;
  q.f0 = c0 >= HALF; q.l0 = q.f0 ? c0 - HALF : c0;
  q.f1 = c1 >= HALF; q.l1 = q.f1 ? c1 - HALF : c1;
  return q;
}

// ---- pass 2: vertical DFT4, Hermitian split, horizontal DFT4.  Task (u, n27'); 16 lanes per n27', 14 used.
B2C_MATH void pass2(int t, cplx* Z) {
  const int u = t & 15;
  if (u >= NU) return;
  const int pl = u ? REGB + u : 0;   // group 0 is its own mirror
#pragma unroll 1
  for (int n = t >> 4; n < 27; n += THREADS / 16) {
    const Pair q = pair_of(n);
    cplx* z0 = Z + q.l0 * LD;
    cplx* z1 = Z + q.l1 * LD;
    cplx own[2][4], par[2][4];
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      own[0][s] = z0[NU * s + u]; par[0][s] = z0[NU * s + pl];
      own[1][s] = z1[NU * s + u]; par[1][s] = z1[NU * s + pl];
    }
    Bfly<4>::run(own[0]); Bfly<4>::run(par[0]); Bfly<4>::run(own[1]); Bfly<4>::run(par[1]);
    cplx col[4][4];   // [horizontal n4'][vertical s]
#pragma unroll
    for (int x = 0; x < 2; ++x) {
      const bool f = x ? q.f1 : q.f0;
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        const cplx a = own[x][s], b = par[x][(4 - s) & 3];
        const cplx re = make_double2(a.x + b.x, a.y - b.y);   // 2 * DFT(real-part column)      at (u, s)
        const cplx im = make_double2(a.y + b.y, b.x - a.x);   // 2 * DFT(imaginary-part column)
        col[x][s] = f ? im : re;
        col[x + 2][s] = f ? re : im;
      }
    }
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      cplx h[4] = {col[0][s], col[1][s], col[2][s], col[3][s]};
      Bfly<4>::run(h);
      z0[NU * s + u] = h[0]; z1[NU * s + u] = h[1];
      z0[REGB + NU * s + u] = h[2]; z1[REGB + NU * s + u] = h[3];
    }
  }
}

// ---- pass 3: horizontal DFT27, kernel spectrum, inverse DFT27.  Thread (k4' = t / 56, kk = 14 s + u = t % 56).
// Lines of the 27 points: (4 n) mod 54 for even k4'.  For odd k4' the points live at (4 n + 27) mod 54 =
// ((4 (n - 7)) mod 54) + 1: the same walk one line further down, started 7 points later.  A cyclic shift of the
// sequence commutes with the product by the spectrum (it is a phase that the inverse removes), so odd threads
// just offset their base by one line and every offset below is a compile-time constant.
B2C_MATH cplx* p3_base(int t, cplx* Z) {
  const int k4 = t / KK, kk = t - k4 * KK;
  return Z + (k4 >> 1) * REGB + kk + ((k4 & 1) ? LD : 0);
}
B2C_MATH void p3_load(cplx* v, const cplx* z) {
#pragma unroll
  for (int n = 0; n < 27; ++n) v[n] = z[((4 * n) % HALF) * LD];
}
// v *= kernel spectrum, swapped for the inverse transform.  The 27 values come from L2 (the spectrum is larger than
// what L1 keeps next to two images' working sets): they are fetched in three volatile groups of nine so that one L2
// latency covers nine of them -- left to itself the compiler sinks every load next to its use (27 exposed latencies,
// measured as the top stall of the kernel).
B2C_MATH cplx ldg_cplx_volatile(const cplx* ptr) {
#ifdef __CUDA_ARCH__
  cplx w;
  asm volatile("ld.global.nc.v2.f64 {%0, %1}, [%2];" : "=d"(w.x), "=d"(w.y) : "l"(ptr));
  return w;
#else
  return *ptr;
#endif
}
B2C_MATH void p3_mul(cplx* v, int t, const cplx* __restrict__ kspec) {
  const cplx* __restrict__ ks = kspec + t / KK * (27 * KK) + t % KK;
#pragma unroll
  for (int g = 0; g < 27; g += 9) {
    cplx w[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) w[k] = ldg_cplx_volatile(ks + (g + k) * KK);
#pragma unroll
    for (int k = 0; k < 9; ++k) v[g + k] = cswap(cmul(v[g + k], w[k]));
  }
}
B2C_MATH void p3_store(const cplx* v, cplx* z) {
#pragma unroll
  for (int n = 0; n < 27; ++n) z[((4 * n) % HALF) * LD] = cswap(v[n]);
}

// ---- pass 4: inverse horizontal DFT4, repack the column pair, inverse vertical DFT4.
B2C_MATH void pass4(int t, cplx* Z) {
  const int u = t & 15;
  if (u >= NU) return;
#pragma unroll 1
  for (int n = t >> 4; n < 27; n += THREADS / 16) {
    const Pair q = pair_of(n);
    cplx* z0 = Z + q.l0 * LD;
    cplx* z1 = Z + q.l1 * LD;
    cplx col[4][4];
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      cplx h[4] = {cswap(z0[NU * s + u]), cswap(z1[NU * s + u]), cswap(z0[REGB + NU * s + u]), cswap(z1[REGB + NU * s + u])};
      Bfly<4>::run(h);
#pragma unroll
      for (int m = 0; m < 4; ++m) col[m][s] = cswap(h[m]);
    }
    cplx own[2][4], par[2][4];
#pragma unroll
    for (int x = 0; x < 2; ++x) {
      const bool f = x ? q.f1 : q.f0;
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        const cplx y1 = f ? col[x + 2][s] : col[x][s], y2 = f ? col[x][s] : col[x + 2][s];
        // swapped on the way into the inverse vertical DFT4
        own[x][s] = make_double2(y1.y + y2.x, y1.x - y2.y);             // Y1 + i*Y2 at (u, s)
        par[x][(4 - s) & 3] = make_double2(y2.x - y1.y, y1.x + y2.y);   // conj(Y1) + i*conj(Y2) at the mirror
      }
    }
    Bfly<4>::run(own[0]); Bfly<4>::run(par[0]); Bfly<4>::run(own[1]); Bfly<4>::run(par[1]);
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      z0[NU * s + u] = cswap(own[0][s]); z1[NU * s + u] = cswap(own[1][s]);
      z0[REGB + NU * s + u] = cswap(par[0][s]); z1[REGB + NU * s + u] = cswap(par[1][s]);   // u == 0: spare slots, never read
    }
  }
}

// frequencies of the kernel-spectrum entry [k4'][k27'][kk = 14 s + u]
__host__ __device__ constexpr int kv_of(int kk) { return (81 * (kk / NU) + 28 * (kk % NU)) % P; }
__host__ __device__ constexpr int kh_of(int k4, int k27) { return (81 * k4 + 28 * k27) % P; }

}  // namespace pfa
}  // namespace b2c
