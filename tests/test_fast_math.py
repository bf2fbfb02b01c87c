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
                '{ for (int i = 0; i < n; ++i) out[i] = vk::fast_log_logistic(x[i]); }\n'
                'extern "C" void fm_eval_tbl(const double* x, int n, double* out)'
                '{ for (int i = 0; i < n; ++i) out[i] = vk::tbl_log_logistic_host(x[i]); }\n' % ROOT)
    so = os.path.join(d, "fm.so")
    subprocess.check_call(["/usr/bin/g++", "-O2", "-fPIC", "-shared", "-mfma", "-o", so, src])
    lib = ctypes.CDLL(so)

    def ev(x, tbl=False):
        x = np.ascontiguousarray(x, dtype=np.float64)
        out = np.empty_like(x)
        (lib.fm_eval_tbl if tbl else lib.fm_eval)(x.ctypes.data_as(ctypes.c_void_p), x.size,
                                                  out.ctypes.data_as(ctypes.c_void_p))
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


def test_table_variant_accuracy_vs_multiprecision(fm):
    """tbl_softplus_neg (the BLR MH hot loop): relative error of log_logistic grows with |x| because a
    single constant is used for ln2/64 (fast_math.cuh); bound checked: 5e-16 + 1.2e-16*|x|."""
    mp.mp.dps = 40
    rng = np.random.default_rng(1)
    x = np.concatenate([rng.uniform(-40, 40, 8000), rng.uniform(-700, 700, 2000), rng.normal(0, 2, 8000),
                        rng.uniform(-1e-3, 1e-3, 500),
                        [0, -37, -36.999, 37, 1e-300, -1e-300, 700, -700, 1e-9, -1e-9, 745, -745, 5000.0, -5000.0]])
    out = fm(x, tbl=True)
    worst = 0.0
    for xi, oi in zip(x, out):
        if abs(xi) > 700:
            continue
        ex = -mp.log1p(mp.exp(-mp.mpf(float(xi))))
        rel = float(abs((mp.mpf(float(oi)) - ex) / ex))
        assert rel < 5e-16 + 1.2e-16 * abs(xi), (xi, oi, rel)
        if abs(xi) <= 40:
            worst = max(worst, rel)
    assert worst < 5e-15, worst
    # beyond the clamp: -|x| for negative x (exact to 1 ulp), ~0 for positive x
    big = fm(np.array([5000.0, -5000.0, 1e308, -1e308]), tbl=True)
    assert abs(big[0]) < 1e-300 and big[1] == -5000.0 and abs(big[2]) < 1e-300 and big[3] == -1e308
    assert np.isnan(fm(np.array([np.nan]), tbl=True)[0])


def test_table_variant_matches_reference_values(fm):
    for c in golden("lite_utils.json")["log_logistic"]:
        x, ref = unj(c["x"]), unj(c["out"])
        if np.isinf(x) or abs(x) > 700:
            continue
        got = fm([x], tbl=True)[0]
        assert abs(got - ref) <= 1e-13 * abs(ref), (x, got, ref)
