"""The FFT-vs-direct choice (auto_path in csrc/capi.cu) against the measured crossover table of BASELINE configs[4]
(profiles/r02_crossover_all_modes.csv: 283 square points, all eight modes, measured on a B200 by
tools/crossover_all_modes.py).  No GPU needed: the choice is a host-side model and the table is committed.
The model may never pick a path that measured more than 15 % slower than the other one."""
import csv
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TABLE = os.path.join(ROOT, "profiles", "r02_crossover_all_modes.csv")


def test_auto_path_within_15_percent_of_the_measured_best(b2c):
    rows = [r for r in csv.reader(open(TABLE)) if r and not r[0].startswith("#")][1:]
    assert len(rows) >= 250
    modes = set()
    worst = (0.0, None)
    for r in rows:
        mode, s, k = int(r[0]), int(r[1]), int(r[2])
        fft, direct = float(r[3]), float(r[4])
        modes.add(mode)
        p = b2c.Plan(mode, s, s, k, k)
        path = p.info()["path"]
        p.destroy()
        assert path in (b2c.PATH_FFT, b2c.PATH_DIRECT)
        chosen = fft if path == b2c.PATH_FFT else direct
        assert chosen > 0, ("the chosen path could not run this point", mode, s, k)
        loss = max(fft, direct) / chosen - 1.0
        if loss > worst[0]:
            worst = (loss, (b2c.MODE_NAMES[mode], s, k, path))
    assert modes == set(range(8))
    assert worst[0] <= 0.15, worst


def test_auto_path_on_the_1d_table(b2c):
    """Same bound on 1 x N rows (profiles/r02_crossover_1d.csv, tools/crossover_1d.py: N = 64 .. 4096, 3 .. 63 taps,
    LINEAR_SAME and LINEAR_FULL), where the bulk-staged one-row direct variant meets the tuned 1-D FFT kernels."""
    table = os.path.join(ROOT, "profiles", "r02_crossover_1d.csv")
    rows = [r for r in csv.reader(open(table)) if r and not r[0].startswith("#")][1:]
    assert len(rows) >= 60
    worst = (0.0, None)
    for r in rows:
        mode, ns, nk = int(r[0]), int(r[1]), int(r[2])
        fft, direct = float(r[3]), float(r[4])
        p = b2c.Plan(mode, 1, ns, 1, nk)
        path = p.info()["path"]
        p.destroy()
        chosen = fft if path == b2c.PATH_FFT else direct
        assert chosen > 0, (mode, ns, nk)
        loss = max(fft, direct) / chosen - 1.0
        if loss > worst[0]:
            worst = (loss, (b2c.MODE_NAMES[mode], ns, nk, path))
    assert worst[0] <= 0.15, worst
