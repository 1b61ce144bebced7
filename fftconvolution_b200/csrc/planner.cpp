// Host-side planner.  See planner.h.  Pure C++ (no CUDA) so that planning works on a CPU-only box.
#include "planner.h"

#include <algorithm>
#include <vector>

namespace b2c {

bool is_smooth2357(int n) {
  if (n < 1) return false;
  for (int p : {2, 3, 5, 7})
    while (n % p == 0) n /= p;
  return n == 1;
}

int next_smooth2357(int n) {
  if (n < 1) n = 1;
  while (!is_smooth2357(n)) ++n;
  return n;
}

// Radices the any-size kernels have a register butterfly for (fft_device.cuh: BigBfly).
// Radix 14 and 16 butterflies exist but spill in the 128-register any-size kernels (measured slower).
static const int kRadices[] = {12, 10, 9, 8, 7, 6, 5, 4, 3, 2};

static double butterfly_cost(int r) {
  switch (r) {
    case 2: return 4;
    case 3: return 12;
    case 4: return 16;
    case 5: return 36;
    case 6: return 36;     // 2 x 3 prime-factor
    case 7: return 66;
    case 8: return 56;
    case 9: return 100;    // 3 x 3 with four internal twiddles
    case 10: return 92;    // 2 x 5 prime-factor
    case 12: return 100;   // 3 x 4 prime-factor
  }
  return 1e9;
}

// Cheapest schedule: fewest passes first (every pass is a trip through shared memory and a barrier),
// then the cheapest butterflies per point.
static void search_schedule(int n, int max_radix, int depth, int* cur, int* best, int* best_len, double cost,
                            double* best_cost) {
  if (n == 1) {
    const double total = cost + 1e6 * depth;
    if (total < *best_cost) {
      *best_cost = total;
      *best_len = depth;
      for (int i = 0; i < depth; ++i) best[i] = cur[i];
    }
    return;
  }
  if (depth >= kMaxStages) return;
  for (int r : kRadices) {
    if (r > max_radix || n % r) continue;   // non-increasing radices: each multiset is visited once
    cur[depth] = r;
    search_schedule(n / r, r, depth + 1, cur, best, best_len, cost + butterfly_cost(r) / r, best_cost);
  }
}

int radix_schedule(int n, int* radix) {
  if (n <= 1) return 0;
  int cur[kMaxStages], best[kMaxStages], len = 0;
  double best_cost = 1e30;
  search_schedule(n, 12, 0, cur, best, &len, 0.0, &best_cost);
  // ascending radices: the last (largest) stage is the one that needs no twiddles
  std::sort(best, best + len);
  for (int i = 0; i < len; ++i) radix[i] = best[i];
  return len;
}

double fft_cost(int n) {
  int radix[kMaxStages];
  int ns = radix_schedule(n, radix);
  double c = 0;
  for (int s = 0; s < ns; ++s) {
    const int r = radix[s];
    const double tw = (s + 1 < ns) ? 4.0 * (r - 1) : 0.0;
    c += (double)(n / r) * (butterfly_cost(r) + tw);
    c += 10.0 * n;  // one load + one store pass over shared memory, index arithmetic
  }
  return c;
}

static std::vector<int> period_candidates(int n_min) {
  std::vector<int> v;
  int p = next_smooth2357(n_min);
  const int limit = n_min + n_min / 4 + 8;
  while ((int)v.size() < 8 && (p <= limit || v.empty())) {
    v.push_back(p);
    p = next_smooth2357(p + 1);
  }
  return v;
}

int choose_period(int n_min) {
  int best = 0;
  double best_c = 0;
  for (int p : period_candidates(n_min)) {
    double c = fft_cost(p);
    if (!best || c < best_c) { best = p; best_c = c; }
  }
  return best;
}

// factorize.h:71-90 -- `n % 4*4*4*2 == 0` is ((n%4)*32)==0 by C precedence: multiples of 4 rejected.
static bool ref_is_optimal(int n) {
  if (n % 4 * 4 * 4 * 2 == 0) return false;
  if (n == 1) return false;
  static const int f[] = {13, 11, 7, 5, 3, 2, 0};
  int ntest = n, last = 0;
  for (int i = 0; f[i] && ntest != 1; ++i)
    while (ntest % f[i] == 0) { ntest /= f[i]; last = f[i]; }
  if (ntest != 1) last = ntest;
  for (int i = 0; f[i]; ++i)
    if (last == f[i]) return true;
  return false;
}

int ref_find_closest_factor(int n) {
  if (ref_is_optimal(n)) return n;
  int j = n + 1;
  while (!ref_is_optimal(j)) ++j;
  return j;
}

static void finish_axis(AxisPlan& a) {
  a.nstages = radix_schedule(a.P, a.radix);
}

// Per-axis facts that do not depend on the chosen period.
static void axis_for_mode(int mode, int ns, int nk, AxisPlan& a) {
  a.n_src = ns;
  a.n_ker = nk;
  const int L = ns + nk - 1;  // full linear length
  a.fold = 0;
  a.kwrap = 0;  // 0 = wrap modulo the period (set after P is known)
  switch (mode) {
    case 0:  // LINEAR_FULL            convolution_fftw.h:54-60
    case 6:  // CIRCULAR_FULL_UNPADDED :106-113
    case 7:  // CIRCULAR_FULL          :114-121  (period == full length: no aliasing)
      a.n_dst = L; a.off = 0; a.min_period = L; a.lin = L; break;
    case 1:  // LINEAR_SAME_UNPADDED   :61-67
    case 2:  // LINEAR_SAME            :68-74 ; crop offset int(hk/2.0) :238-239
      a.n_dst = ns; a.off = nk / 2; a.min_period = ns + nk / 2; a.lin = a.off + ns; break;
    case 3:  // LINEAR_VALID           :75-92 ; crop offset hk-1 :259-260
      a.n_dst = ns - nk + 1; a.off = nk - 1; a.min_period = ns; a.lin = ns; break;
    case 4:  // CIRCULAR_SAME          :93-98 (exact period ns)
      a.n_dst = ns; a.off = 0;
      if (is_smooth2357(ns)) {
        a.exact = true;
        a.min_period = ns; a.lin = ns;  // run at the exact period, kernel wraps modulo ns
      } else {
        // fold(linear result) modulo ns == circular convolution modulo ns
        a.fold = ns; a.kwrap = ns;
        const int nke = std::min(nk, ns);
        a.min_period = ns + nke - 1; a.lin = ns + nke - 1;
      }
      break;
    case 5:  // CIRCULAR_SAME_PADDED   :99-105 == top-left ns crop of the full linear result
      a.n_dst = ns; a.off = 0; a.min_period = L; a.lin = ns; break;
  }
}

// Periods with a tuned shared-kernel kernel: square 2-D (conv_tuned108.cu, conv_tuned2.cu) and 1-D (conv_tuned1d.cu).
static const int kMenu2D[] = {16, 20, 24, 28, 32, 36, 40, 42, 48, 50, 54, 56, 60, 64, 70, 72, 80, 84, 90, 96, 100, 108, 112, 120, 126, 128, 140, 144, 160};
static const int kMenu1D[] = {128, 192, 256, 320, 384, 512, 640, 768, 1024, 1280, 1536, 2048, 2304, 2560, 3072, 4096, 4608, 5120, 6144, 8192};

static bool in_menu(const int* menu, int n, int P) {
  for (int i = 0; i < n; ++i)
    if (menu[i] == P) return true;
  return false;
}

bool plan_on_tuned_menu(const Plan2D& pl) {
  if (pl.empty || pl.h.fold || pl.w.fold || pl.h.kwrap != pl.h.P || pl.w.kwrap != pl.w.P) return false;
  if (pl.h.P == 1 && pl.h.n_src == 1 && pl.h.n_ker == 1) return in_menu(kMenu1D, sizeof kMenu1D / sizeof kMenu1D[0], pl.w.P);
  return pl.h.P == pl.w.P && in_menu(kMenu2D, sizeof kMenu2D / sizeof kMenu2D[0], pl.h.P);
}

// One axis of a CIRCULAR_SAME plan computed as the linear convolution of the periodically extended source (n_in =
// ns + nke - 1 samples, the first nke - 1 a copy of the last ones) cropped at nke - 1.  A kernel longer than the source
// wraps around it (taps congruent modulo ns add up, convolution_fftw.h:185-187): it counts with nke = ns taps and
// kwrap = ns tells the spectrum kernel to fold the taps by ns instead of by the period.
static void ext_axis(AxisPlan& a, int ns, int nk, int P) {
  a = AxisPlan();
  const int nke = std::min(nk, ns);
  a.n_src = ns; a.n_ker = nk; a.n_dst = ns;
  a.off = nke - 1; a.src_shift = nke - 1; a.n_in = ns + nke - 1;
  a.lin = a.min_period = ns + nke - 1;
  a.P = P; a.kwrap = nk > ns ? ns : P; a.n_ker_eff = nke;
  finish_axis(a);
}

bool make_plan_circular_ext(int hs, int ws, int hk, int wk, Plan2D& out) {
  if (hs < 1 || ws < 1 || hk < 1 || wk < 1) return false;
  if (hs == 1 && wk > ws) return false;   // the tuned 1-D spectrum kernel folds by the period only
  out = Plan2D();
  out.mode = 4;
  out.ref_h_fft = hs; out.ref_w_fft = ws;
  if (hs == 1) {   // 1 x N row signal
    const int need = ws + wk - 1;
    for (int pm : kMenu1D) {
      if (pm < need) continue;
      if (pm > 2 * need) return false;
      ext_axis(out.h, 1, 1, 1);
      ext_axis(out.w, ws, wk, pm);
      return true;
    }
    return false;
  }
  if (ws == 1) return false;   // columns are planned as rows by the caller
  const int need = std::max(hs + std::min(hk, hs) - 1, ws + std::min(wk, ws) - 1);
  for (int pm : kMenu2D) {
    if (pm < need) continue;
    ext_axis(out.h, hs, hk, pm);
    ext_axis(out.w, ws, wk, pm);
    return true;
  }
  if (need <= 200) {   // the compile-time 200 x 200 strip kernel (conv_big2ds.cu) reads periodically extended sources too
    ext_axis(out.h, hs, hk, 200);
    ext_axis(out.w, ws, wk, 200);
    return true;
  }
  return false;
}

bool make_plan(int mode, int hs, int ws, int hk, int wk, Plan2D& out) {
  if (mode < 0 || mode > 7) return false;
  out = Plan2D();
  out.mode = mode;
  if (hs < 1 || ws < 1 || hk < 1 || wk < 1) return false;
  // reference sizes, informational
  switch (mode) {
    case 0: case 6: out.ref_h_fft = ref_find_closest_factor(hs + hk - 1); out.ref_w_fft = ref_find_closest_factor(ws + wk - 1); break;
    case 1: out.ref_h_fft = hs + hk / 2; out.ref_w_fft = ws + wk / 2; break;
    case 2: out.ref_h_fft = ref_find_closest_factor(hs + hk / 2); out.ref_w_fft = ref_find_closest_factor(ws + wk / 2); break;
    case 3: out.ref_h_fft = ref_find_closest_factor(hs); out.ref_w_fft = ref_find_closest_factor(ws); break;
    case 4: out.ref_h_fft = hs; out.ref_w_fft = ws; break;
    case 5: out.ref_h_fft = ref_find_closest_factor(hs + hk); out.ref_w_fft = ref_find_closest_factor(ws + wk); break;
    case 7: out.ref_h_fft = hs + hk - 1; out.ref_w_fft = ws + wk - 1; break;
  }
  if (mode == 3 && (hk > hs || wk > ws)) {  // convolution_fftw.h:77-84
    out.empty = true;
    out.ref_h_fft = out.ref_w_fft = 0;
    out.h.n_src = hs; out.h.n_ker = hk; out.w.n_src = ws; out.w.n_ker = wk;
    return true;
  }
  axis_for_mode(mode, hs, hk, out.h);
  axis_for_mode(mode, ws, wk, out.w);

  // Joint choice of (P_h, P_w): row passes touch (n_src + lin) lines of length P_w,
  // column passes touch 2*P_w lines of length P_h.
  std::vector<int> ch = period_candidates(out.h.min_period);
  std::vector<int> cw = period_candidates(out.w.min_period);
  if (out.h.exact) ch.assign(1, out.h.min_period);
  if (out.w.exact) cw.assign(1, out.w.min_period);
  double best = -1;
  for (int ph : ch)
    for (int pw : cw) {
      const double rows = (double)std::min(ph, out.h.n_src) + std::min(ph, out.h.lin);
      double c = rows * fft_cost(pw) + 2.0 * pw * fft_cost(ph) + 12.0 * ph * pw;
      if (ph == 108 && pw == 108) c *= 0.3;  // hand-tuned 108 x 108 kernel (conv_tuned108.cu), measured
      if (best < 0 || c < best) { best = c; out.h.P = ph; out.w.P = pw; }
    }
  // square periods that have a hand-tuned kernel (conv_tuned108.cu, conv_tuned2.cu): one period for
  // both axes, about a third of the generic kernel's time at equal size (measured)
  if (!out.h.exact && !out.w.exact && out.h.fold == 0 && out.w.fold == 0 && hs > 1 && ws > 1) {
    const int need = std::max(out.h.min_period, out.w.min_period);
    for (int pm : kMenu2D) {
      if (pm < need) continue;
      // all tuned kernels share one seven-pass structure: their time scales with the P*P points moved,
      // anchored at 0.3 x the generic model for the measured 108 x 108 case
      static const double alpha = 0.3 * (424.0 * fft_cost(108) + 12.0 * 108 * 108) / (108.0 * 108.0);
      const double c = alpha * pm * pm;
      // (periods below 16 are overhead-bound on the any-size kernel: 10 ns per image at 10 x 10 and 12 x 12 against
      // 2.9 ns for the tuned 16 x 16, profiles/r02_crossover_all_modes.csv)
      if (c < best || pm == 16) { best = std::min(best, c); out.h.P = pm; out.w.P = pm; }
    }
    // beyond the shared-memory kernels: the compile-time 200 x 200 strip kernel (conv_big2ds.cu) takes every period
    // up to 200 -- measured on 100 x 100 (*) k x k LINEAR_FULL: 560 ns per image at 200 against 880-1100 ns for the
    // runtime-schedule strip kernel at 168, 180 and 192 (tools/bench_bigperiods.py)
    // -- i.e. about half the generic model's cost at equal period; rectangles far from square keep their own periods
    if (need > kMenu2D[sizeof kMenu2D / sizeof kMenu2D[0] - 1] && need <= 200) {
      const double rows = (double)std::min(200, out.h.n_src) + std::min(200, out.h.lin);
      const double c = 0.5 * (rows * fft_cost(200) + 2.0 * 200 * fft_cost(200) + 12.0 * 200 * 200);
      if (c < best) { best = c; out.h.P = 200; out.w.P = 200; }
    }
  }
  // 1-D row signals: periods with a tuned kernel (conv_tuned1d.cu: three or four big-radix passes instead of
  // the any-size kernel's five or more small ones, ~3x faster per point) are worth up to 60 % more points
  if (hs == 1 && hk == 1 && !out.w.exact && out.w.fold == 0) {
    for (int pm : kMenu1D) {
      if (pm < out.w.min_period) continue;
      // (short signals: even twice the points beats the any-size kernel, which is overhead-bound there)
      if (pm <= out.w.min_period + (out.w.min_period * 3) / 5 || (pm <= 512 && pm <= 2 * out.w.min_period)) out.w.P = pm;
      break;
    }
  }
  for (AxisPlan* a : {&out.h, &out.w}) {
    if (a->kwrap == 0) a->kwrap = a->P;
    a->n_ker_eff = std::min(a->n_ker, a->kwrap);
    a->lin = std::min(a->lin, a->P);
    a->n_in = a->n_src;
    finish_axis(*a);
  }
  return true;
}

void dif_permutation(int n, int nstages, const int* radix, int* freq_of_pos) {
  for (int pos = 0; pos < n; ++pos) {
    int rem = pos, len = n, mult = 1, freq = 0;
    for (int s = 0; s < nstages; ++s) {
      const int m = len / radix[s];
      const int p = rem / m;
      rem -= p * m;
      freq += p * mult;
      mult *= radix[s];
      len = m;
    }
    freq_of_pos[pos] = freq;
  }
}

}  // namespace b2c
