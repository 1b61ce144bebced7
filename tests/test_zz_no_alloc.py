"""GPU test of b2c_convolve_batch_device_static, the entry point for callers that cannot allocate
device memory (its scratch is a fixed device array inside the library, see include/b2conv.h).

It is called through the C-ABI and compared with the regular entry point
b2c_convolve_batch_device on the same inputs and with the CPU oracle (long-double nested loops) on
the first and last image of every case, within the 1e-12 relative tolerance the other parity tests
use -- twice in a row with different kernels, so a spectrum left in the shared scratch by the first
call would be caught.  A third test makes the very first call of a fresh process on a non-blocking
stream, which is not ordered after the set-up's blocking copy unless the library waits for it.  The file is named so that pytest collects it last: a
run that stops at the first failure has then already run every other test.  It does not depend
on state left by other tests and also passes alone in a fresh process (the first call performs the
per-device set-up itself)."""
import ctypes
import os
import subprocess
import sys

import numpy as np
import pytest

import _oracle

pytestmark = pytest.mark.gpu

REL_TOL = 1e-12


def _rel(a, b):
    return float(np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-300))


def _static_call(b2c, torch, mode, src, ker):
    n, hs, ws = src.shape
    hk, wk = ker.shape
    plan = b2c.Plan(mode, hs, ws, hk, wk)          # only used for the output shape and the comparison
    d_src, d_ker = torch.from_numpy(src).cuda(), torch.from_numpy(ker).cuda()
    d_new = torch.full((n, plan.h_dst, plan.w_dst), float("nan"), dtype=torch.float64, device="cuda")
    d_old = torch.full((n, plan.h_dst, plan.w_dst), float("nan"), dtype=torch.float64, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    rc = b2c.lib().b2c_convolve_batch_device_static(mode, hs, ws, hk, wk, d_src.data_ptr(), d_ker.data_ptr(),
                                                    d_new.data_ptr(), n, ctypes.c_void_p(stream))
    assert rc == 0, b2c.lib().b2c_last_error()
    plan.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), 1, d_old.data_ptr(), n, stream)
    torch.cuda.synchronize()
    plan.destroy()
    return d_new.cpu().numpy(), d_old.cpu().numpy()


def _oracle_check(oracle, mode, src, ker, new, what):
    for i in (0, src.shape[0] - 1):
        want = oracle.convolve_direct8(mode, src[i], ker)
        assert new[i].shape == want.shape, what
        assert _oracle.rel_err(new[i], want) <= REL_TOL, (what, i, _oracle.rel_err(new[i], want))


def test_static_scratch_entry_matches_regular_entry_twice_in_a_row(b2c, oracle):
    import torch
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    rng = np.random.default_rng(99)
    # two calls in a row on the headline shape, different kernels: the second must not see the first's spectrum
    src = rng.random((9, 100, 100))
    for call in range(2):
        ker = rng.random((15, 15)) * (call + 1)
        new, old = _static_call(b2c, torch, b2c.LINEAR_SAME, src, ker)
        assert np.all(np.isfinite(new))
        assert _rel(new, old) <= REL_TOL, ("call %d" % call, _rel(new, old))
        _oracle_check(oracle, b2c.LINEAR_SAME, src, ker, new, "call %d" % call)
    # other modes / shapes that fit the 108 x 108 limit (run at the fixed period 108)
    for mode, hs, ws, hk, wk in [(b2c.LINEAR_FULL, 50, 60, 30, 20), (b2c.LINEAR_VALID, 90, 108, 7, 9),
                                 (b2c.CIRCULAR_SAME_PADDED, 64, 64, 10, 12), (b2c.CIRCULAR_FULL, 1, 80, 1, 29),
                                 (b2c.CIRCULAR_SAME, 108, 108, 20, 130), (b2c.LINEAR_SAME_UNPADDED, 33, 1, 8, 1)]:
        s, k = rng.random((3, hs, ws)), rng.random((hk, wk))
        new, old = _static_call(b2c, torch, mode, s, k)
        assert new.shape == old.shape
        assert _rel(new, old) <= REL_TOL, (mode, hs, ws, hk, wk, _rel(new, old))
        _oracle_check(oracle, mode, s, k, new, (mode, hs, ws, hk, wk))


def test_static_scratch_entry_rejects_what_does_not_fit(b2c):
    import torch
    torch.cuda.set_device(0)
    L = b2c.lib()
    d = torch.zeros(200 * 200, dtype=torch.float64, device="cuda")
    # LINEAR_SAME 100 + 17/2 = 108 fits, 100 + 19/2 = 109 does not
    rc = L.b2c_convolve_batch_device_static(b2c.LINEAR_SAME, 100, 100, 19, 19, d.data_ptr(), d.data_ptr(), d.data_ptr(), 1, None)
    assert rc == 4 and b"108" in L.b2c_last_error()
    rc = L.b2c_convolve_batch_device_static(b2c.CIRCULAR_SAME, 100, 100, 5, 5, d.data_ptr(), d.data_ptr(), d.data_ptr(), 1, None)
    assert rc == 4 and b"CIRCULAR_SAME" in L.b2c_last_error()
    rc = L.b2c_convolve_batch_device_static(12, 10, 10, 3, 3, d.data_ptr(), d.data_ptr(), d.data_ptr(), 1, None)
    assert rc == 1
    # empty LINEAR_VALID is a no-op, like the regular path
    assert L.b2c_convolve_batch_device_static(b2c.LINEAR_VALID, 4, 4, 5, 5, d.data_ptr(), d.data_ptr(), d.data_ptr(), 1, None) == 0
    torch.cuda.synchronize()


_FIRST_CALL = r"""
import ctypes, sys
import numpy as np, torch
sys.path.insert(0, %r); sys.path.insert(0, %r)
import _oracle, fftconvolution_b200 as b2c
torch.cuda.set_device(0)
rng = np.random.default_rng(5)
src, ker = rng.random((4, 100, 100)), rng.random((15, 15))
d_src, d_ker = torch.from_numpy(src).cuda(), torch.from_numpy(ker).cuda()
d_dst = torch.full((4, 100, 100), float("nan"), dtype=torch.float64, device="cuda")
torch.cuda.synchronize()
st = torch.cuda.Stream()          # torch creates its streams non-blocking: no implicit order with the legacy stream
rc = b2c.lib().b2c_convolve_batch_device_static(b2c.LINEAR_SAME, 100, 100, 15, 15, d_src.data_ptr(), d_ker.data_ptr(),
                                                d_dst.data_ptr(), 4, ctypes.c_void_p(st.cuda_stream))
assert rc == 0, b2c.lib().b2c_last_error()
st.synchronize()
got = d_dst.cpu().numpy()
orc = _oracle.load_oracle()
err = max(_oracle.rel_err(got[i], orc.convolve_direct8(b2c.LINEAR_SAME, src[i], ker)) for i in (0, 3))
assert err <= 1e-12, err
print("first-call-ok", err)
"""


def test_first_call_of_a_process_on_a_non_blocking_stream(b2c):
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, "-c", _FIRST_CALL % (root, os.path.join(root, "tests"))], capture_output=True,
                         text=True, timeout=600)
    assert out.returncode == 0 and "first-call-ok" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]
