"""Randomised parity sweep: random shapes x all modes x {auto, fft, generic, direct} x {shared, per-image}
x {batch, single call} against the oracle (long-double nested loops).  Exploration tool, not part of the
test suite; prints every failure and a summary.  usage: python tools/fuzz_parity.py [cases] [seed]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch, fftconvolution_b200 as b2c, _oracle
ncases = int(sys.argv[1]) if len(sys.argv) > 1 else 300
seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
rng = np.random.default_rng(seed)
orc = _oracle.load_oracle()
torch.cuda.set_device(0)
fails = 0; ran = 0
def pick_dim(big):
    r = rng.random()
    if r < 0.15: return 1
    if r < 0.6: return int(rng.integers(2, 40))
    return int(rng.integers(40, big))
for case in range(ncases):
    one_d = rng.random() < 0.25
    if one_d:
        hs, hk = 1, 1
        ws = int(rng.integers(1, 5000)) if rng.random() < 0.5 else int(rng.integers(1, 300))
        wk = int(rng.integers(1, min(600, 2 * ws + 2)))
        if rng.random() < 0.3: hs, ws, hk, wk = ws, hs, wk, hk
    else:
        big = 260 if rng.random() < 0.15 else 140
        hs, ws = pick_dim(big), pick_dim(big)
        hk, wk = pick_dim(60), pick_dim(60)
    mode = int(rng.integers(0, 8))
    shared = int(rng.random() < 0.7)
    n = int(rng.choice([1, 2, 3, 7, 40]))
    if hs * ws * hk * wk * n > 3e8: n = 1
    if hs * ws * hk * wk > 4e8: continue
    try:
        p = b2c.Plan(mode, hs, ws, hk, wk)
    except b2c.B2CError as e:
        print("plan failed", mode, hs, ws, hk, wk, e); fails += 1; continue
    if p.h_dst == 0:
        p.destroy(); continue
    # magnitudes of source and kernel vary independently over twelve decades, images of a batch over eight
    s_scale, k_scale = 10.0 ** rng.uniform(-6, 6), 10.0 ** rng.uniform(-6, 6)
    src = s_scale * (rng.random((n, hs, ws)) - 0.3) * (10.0 ** rng.uniform(-4, 4, size=(n, 1, 1)));   # neighbours differ too
    ker = k_scale * ((rng.random((hk, wk)) if shared else rng.random((n, hk, wk))) - 0.3)
    want = [orc.convolve_direct8(mode, src[i], ker if shared else ker[i]) for i in sorted({0, n - 1})]
    d_src, d_ker = torch.from_numpy(src).cuda(), torch.from_numpy(ker).cuda()
    for path in (b2c.PATH_AUTO, b2c.PATH_FFT, b2c.PATH_FFT_GENERIC, b2c.PATH_DIRECT):
        try:
            p.set_path(path)
            d_dst = torch.full((n, p.h_dst, p.w_dst), float("nan"), dtype=torch.float64, device="cuda")
            p.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), shared, d_dst.data_ptr(), n, 0)
            torch.cuda.synchronize()
        except b2c.B2CError as e:
            if e.code == 4: continue      # documented size limits of a forced path
            print("call failed", (mode, hs, ws, hk, wk, n, shared, path), e); fails += 1; continue
        got = d_dst.cpu().numpy()
        ran += 1
        for w, i in zip(want, sorted({0, n - 1})):
            err = _oracle.rel_err(got[i], w)
            if not (err <= 1e-12) or got[i].shape != w.shape:
                print("MISMATCH", (mode, hs, ws, hk, wk, n, shared, path, i), "rel err", err); fails += 1
        if path == b2c.PATH_AUTO and rng.random() < 0.3:
            one = p.convolve_host(src[0], ker if shared else ker[0])
            err = _oracle.rel_err(one, want[0])
            if not (err <= 1e-12): print("MISMATCH single call", (mode, hs, ws, hk, wk), err); fails += 1
    p.destroy()
print("fuzz: %d kernel runs over %d cases, %d failures (seed %d)" % (ran, ncases, fails, seed))
sys.exit(1 if fails else 0)
