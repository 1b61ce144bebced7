"""Which period the new planner gives a problem decides which kernel runs it (csrc/planner.cpp: make_plan).  These are
the routing rules round 2 added on top of the cost model, pinned on the shapes they were measured on (DESIGN.md 3.2a,
3.2b): the tuned square menu reaches 160, square-ish problems whose period falls in 161..200 take the compile-time
200 x 200 strip kernel, needs below 16 take the tuned 16 x 16 kernel, rectangles far from square keep their own periods.
No GPU needed: planning is host code."""
import pytest


@pytest.mark.parametrize("mode,hs,ws,hk,wk,period", [
    ("LINEAR_SAME", 100, 100, 15, 15, (108, 108)),      # the headline shape: needs 107
    ("LINEAR_FULL", 100, 100, 30, 30, (140, 140)),      # 129 -> menu extension
    ("LINEAR_FULL", 100, 100, 45, 45, (144, 144)),
    ("LINEAR_FULL", 100, 100, 61, 61, (160, 160)),
    ("LINEAR_FULL", 100, 100, 62, 62, (200, 200)),      # 161: beyond the shared-memory kernels
    ("LINEAR_FULL", 100, 100, 90, 90, (200, 200)),
    ("CIRCULAR_FULL", 100, 100, 100, 100, (200, 200)),  # BASELINE configs[2]
    ("LINEAR_FULL", 120, 90, 70, 100, (200, 200)),      # square-ish rectangle: one square period
    ("LINEAR_FULL", 10, 10, 3, 3, (16, 16)),            # tiny: overhead-bound on the any-size kernel
    ("LINEAR_VALID", 12, 12, 3, 3, (16, 16)),
])
def test_square_problems_land_on_the_tuned_periods(b2c, mode, hs, ws, hk, wk, period):
    p = b2c.Plan(getattr(b2c, mode), hs, ws, hk, wk)
    try:
        assert (p.h_fft, p.w_fft) == period
    finally:
        p.destroy()


def test_rectangles_far_from_square_keep_their_own_periods(b2c):
    for hs, ws, hk, wk in [(150, 20, 20, 10), (100, 60, 70, 30)]:
        p = b2c.Plan(b2c.LINEAR_FULL, hs, ws, hk, wk)
        try:
            assert p.h_fft >= hs + hk - 1 and p.w_fft >= ws + wk - 1
            assert p.w_fft < 128 and (p.h_fft, p.w_fft) != (200, 200), (hs, ws, hk, wk, p.h_fft, p.w_fft)
        finally:
            p.destroy()
