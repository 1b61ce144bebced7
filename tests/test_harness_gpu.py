"""The reference's OWN drivers, byte-unchanged, running on the drop-in header + libb2conv.so
(oracle/harness_fftw.cc, oracle/harness_bench.cc explain how).  9 size categories x 50 random
cases x 6 modes, tolerance 1e-10 absolute L2 (convolution_test.cc:49), Octave replaced by the CPU
oracle shim (Octave is not installed on this image)."""
import os
import re
import subprocess

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TEST_BIN = os.path.join(ROOT, "oracle", "_ref", "convolution_test_b2c.bin")
BENCH_BIN = os.path.join(ROOT, "oracle", "_ref", "convolution_benchmark_b2c.bin")


GSL_TEST_BIN = os.path.join(ROOT, "oracle", "_ref", "convolution_test_gsl_b2c.bin")


@pytest.mark.parametrize("flavour", ["fftw", "gsl"])
def test_reference_test_driver_passes_unchanged(tmp_path, flavour):
    """-DCONVOLUTION=0 against include/convolution_fftw.h and -DCONVOLUTION=1 against the alias header
    include/convolution_gsl.h (the reference's two FFT back ends share one interface)."""
    binary = TEST_BIN if flavour == "fftw" else GSL_TEST_BIN
    if not os.path.exists(binary):
        pytest.skip("%s was not built (reference sources absent at build time)" % binary)
    r = subprocess.run([binary], cwd=tmp_path, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=600)
    out = r.stdout.decode("utf-8", "replace")
    assert r.returncode == 0, out[-2000:]
    passed = re.findall(r"All tests passed for ([a-z_ ]+?) (\d+) / (\d+)", out)
    failed = re.findall(r"Some tests failed", out)
    assert not failed, out[-3000:]
    # 9 categories x {full, same, valid, circular full, same unpadded, circular full unpadded}
    assert len(passed) == 54, (len(passed), out[-2000:])
    assert all(a == b == "50" for _, a, b in passed)
    log = open(os.path.join(tmp_path, "convolution_test_%s.log" % flavour)).read()
    assert log.count("PASSED") == 54 * 50 and "FAILED" not in log


def test_reference_benchmark_driver_runs_unchanged(tmp_path):
    """BASELINE configs[0]: 50x50 (*) 10x10 LINEAR_SAME through convolve(), timed by the reference's
    own benchmark main (mean of 20 calls)."""
    if not os.path.exists(BENCH_BIN):
        pytest.skip("oracle/_ref/convolution_benchmark_b2c.bin was not built")
    r = subprocess.run([BENCH_BIN, "50", "10"], cwd=tmp_path, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=120)
    out = r.stdout.decode("utf-8", "replace")
    assert r.returncode == 0, out
    m = re.search(r"([0-9.]+e[-+][0-9]+) s\.", out)
    assert m, out
    # mean of 20 calls INCLUDING the first one (CUDA context + module load of a fresh process, up to
    # ~1 s on a cold box): a sanity bound, not a performance claim -- bench.py measures latency
    assert 0.0 < float(m.group(1)) < 2.0, out
