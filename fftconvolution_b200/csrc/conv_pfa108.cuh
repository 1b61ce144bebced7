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

constexpr int P = 108, HALF = 54, LD = 113, NU = 14, KK = 56, REGB = 56, THREADS = 256;
constexpr size_t SMEM_BYTES = sizeof(cplx) * (size_t)HALF * LD;   // 97 632: two CTAs per SM inside the 196 KB carve-out

#ifdef __CUDA_ARCH__
#define B2C_PFA_LDG(ptr) __ldg(ptr)
#else
#define B2C_PFA_LDG(ptr) (*(ptr))
#endif

// exp(-2 pi i m / 27) for the exponents the 3 x 9 decomposition needs
__host__ __device__ constexpr double w27_re(int m) {
  return m == 1 ? 0.9730448705798238388329 : m == 2 ? 0.8936326403234122481926 : m == 3 ? 0.7660444431189780352024
       : m == 4 ? 0.5971585917027861648519 : m == 5 ? 0.396079766039156823696 : m == 6 ? 0.1736481776669303488517
       : m == 7 ? -0.05814482891047582853875 : m == 8 ? -0.2868032327110902531033 : m == 10 ? -0.6862416378687335857296
       : m == 12 ? -0.9396926207859083840541 : m == 14 ? -0.9932383577419429885479 : m == 16 ? -0.8354878114129364196538 : 1.0;
}
__host__ __device__ constexpr double w27_im(int m) {
  return m == 1 ? -0.2306158707424401784502 : m == 2 ? -0.448799180200462172785 : m == 3 ? -0.6427876096865393263226
       : m == 4 ? -0.8021231927550437850833 : m == 5 ? -0.918216106880274014759 : m == 6 ? -0.9848077530122080593667
       : m == 7 ? -0.9983081582712682080478 : m == 8 ? -0.9579895123154888744374 : m == 10 ? -0.7273736415730486959872
       : m == 12 ? -0.3420201433256687330441 : m == 14 ? 0.1160929141252302296757 : m == 16 ? 0.5495089780708060352628 : 0.0;
}

// Forward 27-point DFT in registers, natural order in and out: 27 = 3 x 9 Cooley-Tukey,
// n = 9 a + b, k = ka + 3 kb:  DFT3 over a, twiddle W27^(b ka), DFT9 over b.  436 FP64 instructions.
B2C_MATH void dft27(cplx* v) {
#pragma unroll
  for (int b = 0; b < 9; ++b) {
    cplx t[3] = {v[b], v[b + 9], v[b + 18]};
    Bfly<3>::run(t);
    v[b] = t[0];
    v[b + 9] = b ? cmul(t[1], make_double2(w27_re(b), w27_im(b))) : t[1];
    v[b + 18] = b ? cmul(t[2], make_double2(w27_re(2 * b), w27_im(2 * b))) : t[2];
  }
  cplx o[27];
#pragma unroll
  for (int ka = 0; ka < 3; ++ka) {
    cplx s[9];
#pragma unroll
    for (int b = 0; b < 9; ++b) s[b] = v[9 * ka + b];
    BigBfly<9>::run(s);
#pragma unroll
    for (int kb = 0; kb < 9; ++kb) o[ka + 3 * kb] = s[kb];
  }
#pragma unroll
  for (int k = 0; k < 27; ++k) v[k] = o[k];
}

// position of vertical (group g = k27, 4-point index s) inside a line
__host__ __device__ constexpr int vpos(int g, int s) { return g < NU ? NU * s + g : REGB + NU * s + (27 - g); }

// Rows of thread J = n4 in ascending order: row RHO + 4 m is element n27 = (m - O) mod 27 of its 27-point
// sequence ((27 J + 4 n27) mod 108).  J is a template parameter (and warp-uniform at run time), so every row
// number, register index and -- for stores into Z -- offset below is a compile-time constant.
template <int J>
struct Rows {
  static constexpr int RHO = (27 * J) & 3, O = (27 * J - RHO) / 4;
  static __host__ __device__ constexpr int n_of(int m) { return (m - O + 27) % 27; }
};

template <int J>
B2C_MATH void load_rows(cplx* v, const Params& p, const double* __restrict__ x, int line) {
  typedef Rows<J> R;
  const bool c0 = line < p.ws_in, c1 = line + HALF < p.ws_in;
  if ((p.sh_h | p.sh_w) == 0) {
    // rows RHO + 4 m < hs  <=>  m < mv; the column tests are folded into the two row limits
    const int mv = (p.hs - R::RHO + 3) >> 2;
    const int mv0 = c0 ? mv : 0, mv1 = c1 ? mv : 0;
    const unsigned ws = (unsigned)p.ws, base = (unsigned)line + R::RHO * ws;
#pragma unroll
    for (int m = 0; m < 27; ++m) {
      const unsigned off = base + (4u * m) * ws;
      v[R::n_of(m)].x = m < mv0 ? B2C_PFA_LDG(x + off) : 0.0;
      v[R::n_of(m)].y = m < mv1 ? B2C_PFA_LDG(x + off + HALF) : 0.0;
    }
  } else {   // source extended periodically by (sh_h, sh_w) samples in front
    int sc0 = line - p.sh_w, sc1 = line + HALF - p.sh_w;
    if (sc0 < 0) sc0 += p.ws;
    if (sc1 < 0) sc1 += p.ws;
#pragma unroll
    for (int m = 0; m < 27; ++m) {
      const int r = R::RHO + 4 * m;
      int sr = r - p.sh_h;
      if (sr < 0) sr += p.hs;
      const bool in = r < p.hs_in;
      v[R::n_of(m)].x = (in && c0) ? B2C_PFA_LDG(x + (unsigned)(sr * p.ws + sc0)) : 0.0;
      v[R::n_of(m)].y = (in && c1) ? B2C_PFA_LDG(x + (unsigned)(sr * p.ws + sc1)) : 0.0;
    }
  }
}

template <int J>
B2C_MATH void store_rows(const cplx* v, const Params& p, double* __restrict__ y, int line) {
  typedef Rows<J> R;
  const int j0 = line - p.off_w, j1 = line + HALF - p.off_w;
  // output row i = RHO + 4 m - off_h must lie in [0, hd); the column tests are folded into the two limits
  const unsigned hd0 = (j0 >= 0 && j0 < p.wd) ? (unsigned)p.hd : 0u, hd1 = (j1 >= 0 && j1 < p.wd) ? (unsigned)p.hd : 0u;
  const unsigned wd = (unsigned)p.wd;
  const unsigned base = (unsigned)((R::RHO - p.off_h) * p.wd + j0);   // wraps for rows above the window: never used then
#pragma unroll
  for (int m = 0; m < 27; ++m) {
    const unsigned i = (unsigned)(R::RHO + 4 * m - p.off_h);
    const unsigned off = base + (4u * m) * wd;
    // v is still re/im swapped: .y is column `line`, .x is column `line + 54`
    if (i < hd0) y[off] = v[R::n_of(m)].y;
    if (i < hd1) y[off + HALF] = v[R::n_of(m)].x;
  }
}

// ---- pass 1: vertical 27-point DFTs, global -> Z.  Thread (n4 = j = t / 64, line = t % 64 < 54).
B2C_MATH void pass1(int t, cplx* Z, const Params& p, const double* __restrict__ x) {
  const int j = t >> 6, line = t & 63;
  if (line >= HALF) return;
  cplx v[27];
  switch (j) {
    case 0: load_rows<0>(v, p, x, line); break;
    case 1: load_rows<1>(v, p, x, line); break;
    case 2: load_rows<2>(v, p, x, line); break;
    default: load_rows<3>(v, p, x, line); break;
  }
  dft27(v);
  cplx* z = Z + line * LD + NU * j;
#pragma unroll
  for (int k = 0; k < 27; ++k) z[vpos(k, 0)] = v[k];
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

// ---- pass 3: horizontal DFT27, kernel spectrum, inverse DFT27.  Thread (k4', kk = 14 s + u), 224 threads.
// Lines of the 27 points: (4 n) mod 54 for even k4'.  For odd k4' the points live at (4 n + 27) mod 54 =
// ((4 (n - 7)) mod 54) + 1: the same walk one line further down, started 7 points later.  A cyclic shift of the
// sequence commutes with the product by the spectrum (it is a phase that the inverse removes), so odd threads
// just offset their base by one line and every offset below is a compile-time constant.
B2C_MATH void pass3(int t, cplx* Z, const cplx* __restrict__ kspec) {
  if (t >= 4 * KK) return;
  const int k4 = t / KK, kk = t - k4 * KK;
  cplx* z = Z + (k4 >> 1) * REGB + kk + ((k4 & 1) ? LD : 0);
  cplx v[27];
#pragma unroll
  for (int n = 0; n < 27; ++n) v[n] = z[((4 * n) % HALF) * LD];
  dft27(v);
  const cplx* __restrict__ ks = kspec + (k4 * 27) * KK + kk;
#pragma unroll
  for (int k = 0; k < 27; ++k) v[k] = cswap(cmul(v[k], B2C_PFA_LDG(ks + k * KK)));   // swapped for the inverse
  dft27(v);
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

// ---- pass 5: inverse vertical DFT27, crop, Z -> global.
B2C_MATH void pass5(int t, const cplx* Z, const Params& p, double* __restrict__ y) {
  const int j = t >> 6, line = t & 63;
  if (line >= HALF) return;
  const cplx* z = Z + line * LD + NU * j;
  cplx v[27];
#pragma unroll
  for (int k = 0; k < 27; ++k) v[k] = cswap(z[vpos(k, 0)]);
  dft27(v);
  switch (j) {
    case 0: store_rows<0>(v, p, y, line); break;
    case 1: store_rows<1>(v, p, y, line); break;
    case 2: store_rows<2>(v, p, y, line); break;
    default: store_rows<3>(v, p, y, line); break;
  }
}

// frequencies of the kernel-spectrum entry [k4'][k27'][kk = 14 s + u]
__host__ __device__ constexpr int kv_of(int kk) { return (81 * (kk / NU) + 28 * (kk % NU)) % P; }
__host__ __device__ constexpr int kh_of(int k4, int k27) { return (81 * k4 + 28 * k27) % P; }

}  // namespace pfa
}  // namespace b2c
