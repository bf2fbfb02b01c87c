"""CPU: the hand-scheduled fp64 log_logistic of the logistic-regression kernels
(venturecxx_b200/csrc/fast_math.cuh), compiled for the host from the same header,
against a 40-digit reference and the reference's own log_logistic values."""
import ctypes
import os
import subprocess
import tempfile

import mpmath as mp
import numpy as np
import pytest

from helpers import golden, unj

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def fm():
    d = tempfile.mkdtemp()
    src = os.path.join(d, "fm.cpp")
    with open(src, "w") as f:
        f.write('#define VK_HD inline\n#define __host__\n#define __device__\n'
                '#include "%s/venturecxx_b200/csrc/fast_math.cuh"\n'
                'extern "C" void fm_eval(const double* x, int n, double* out)'
                '{ for (int i = 0; i < n; ++i) out[i] = vk::fast_log_logistic(x[i]); }\n' % ROOT)
    so = os.path.join(d, "fm.so")
    subprocess.check_call(["/usr/bin/g++", "-O2", "-fPIC", "-shared", "-mfma", "-o", so, src])
    lib = ctypes.CDLL(so)

    def ev(x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        out = np.empty_like(x)
        lib.fm_eval(x.ctypes.data_as(ctypes.c_void_p), x.size, out.ctypes.data_as(ctypes.c_void_p))
        return out
    return ev


def test_relative_accuracy_vs_multiprecision(fm):
    mp.mp.dps = 40
    rng = np.random.default_rng(0)
    x = np.concatenate([rng.uniform(-40, 40, 6000), rng.uniform(-700, 700, 1500), rng.normal(0, 2, 6000),
                        [0, -37, -36.999, 37, 1e-300, -1e-300, 700, -700]])
    out = fm(x)
    worst = 0.0
    for xi, oi in zip(x, out):
        ex = -mp.log1p(mp.exp(-mp.mpf(float(xi))))
        worst = max(worst, float(abs((mp.mpf(float(oi)) - ex) / ex)))
    assert worst < 6e-16, worst


def test_matches_reference_log_logistic_values(fm):
    # values produced by the reference's utils.log_logistic (tests/golden/make_golden.py)
    for c in golden("lite_utils.json")["log_logistic"]:
        x, ref = unj(c["x"]), unj(c["out"])
        got = fm([x])[0]
        if np.isinf(x):
            # -inf -> -inf like the reference; +inf is outside the kernel's domain (inf - inf)
            assert (got == ref) if x < 0 else (abs(got) < 1e-300 or np.isnan(got))
        elif abs(x) > 700:
            assert abs(got - ref) <= 1e-300 + 1e-15 * abs(ref)
        else:
            assert abs(got - ref) <= 1e-15 * abs(ref), (x, got, ref)
