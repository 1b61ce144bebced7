// Host-side size/schedule planner of the B200 convolution engine.
// Replaces find_closest_factor (reference Convolution/src/factorize.h:92-104) and the mode switch
// of init_workspace (Convolution/src/convolution_fftw.h:52-132).
#pragma once

namespace b2c {

constexpr int kMaxStages = 16;

// Every convolution mode is "linear convolution at an alias-safe period P, then crop at `off`
// or fold modulo `fold`" (SURVEY.md 8a-8).  One AxisPlan per axis.
struct AxisPlan {
  int n_src = 0, n_ker = 0;
  int n_dst = 0;      // output extent, identical to the reference's h_dst / w_dst
  int off = 0;        // first index of the linear result that is kept
  bool exact = false; // period must equal min_period (CIRCULAR_SAME run at its own smooth size)
  int fold = 0;       // >0: output = linear result folded modulo `fold` (CIRCULAR_SAME, non-smooth size)
  int kwrap = 0;      // kernel taps are wrap-accumulated modulo kwrap before anything else (>= 1)
  int n_ker_eff = 0;  // min(n_ker, kwrap): extent of the wrapped kernel
  int lin = 0;        // number of leading entries of the (periodic) linear result that are needed
  int min_period = 0; // smallest alias-safe period
  int P = 0;          // chosen 2/3/5/7-smooth period (>= min_period)
  // Periodic-extension form of CIRCULAR_SAME (make_plan_circular_ext): the transform input has n_in = n_src +
  // n_ker - 1 positions, position q holding source sample (q - src_shift) mod n_src; 0 / n_src otherwise
  int src_shift = 0;
  int n_in = 0;
  int nstages = 0;
  int radix[kMaxStages] = {0};
};

struct Plan2D {
  int mode = 0;
  bool empty = false;  // LINEAR_VALID with kernel > source: 0x0 result, convolve is a no-op
  AxisPlan h, w;
  int ref_h_fft = 0, ref_w_fft = 0;  // what the reference planner would have used (informational)
};

bool is_smooth2357(int n);
int next_smooth2357(int n);
// Factor a 2/3/5/7-smooth n into the radix schedule (radices from {2,3,4,5,7,8}, ascending).
int radix_schedule(int n, int* radix);
// Cost model (FP64 instructions + shared-memory passes, arbitrary units) of one length-n FFT.
double fft_cost(int n);
// Cheapest smooth period >= n_min for one axis, `lines_weight` = relative weight of this axis.
int choose_period(int n_min);

// Reference planner restated (factorize.h:71-104, FFTW_FACTORS = {13,11,7,5,3,2}).
int ref_find_closest_factor(int n);

// Returns false for an unknown mode.
bool make_plan(int mode, int hs, int ws, int hk, int wk, Plan2D& out);

// CIRCULAR_SAME as a plain linear problem: y[d] = sum_j k[j] x[(d - j) mod n] is the linear convolution of the
// source extended periodically by n_ker - 1 samples in front, cropped at n_ker - 1 -- alias-safe at ANY period
// >= n_src + n_ker - 1, so the tuned kernels (menu periods) can run sizes whose own period is not on a menu
// or not even smooth.  Returns true (and the plan) only when kernel <= source on both axes and the periods
// land on a tuned menu (square 2-D menu or 1-D menu).
bool make_plan_circular_ext(int hs, int ws, int hk, int wk, Plan2D& out);
// True when the periods of a plan are served by a tuned shared-kernel kernel (square / 1-D menu, no folding).
bool plan_on_tuned_menu(const Plan2D& pl);

// Position <-> frequency maps of the in-place decimation-in-frequency transform with the
// given schedule: freq_of_pos[p] is the frequency index stored at position p after the forward
// transform.  Arrays of length n.
void dif_permutation(int n, int nstages, const int* radix, int* freq_of_pos);

}  // namespace b2c
