"""GPU parity for config C5: subsampled (austerity) MH on the logistic-regression
model vs the oracle, which evaluates every local section and runs the reference's
sequentialTest (infer/subsampled_mh.py:76-171, 377-409), plus the posterior check
of test/inference_language/test_subsampledmh.py:64-97 in spirit (epsilon -> exact MH)."""
import numpy as np
import pytest

import oracle
import venturecxx_b200 as vb
from test_gpu_blr import make_data, orc_loglik, read_w

pytestmark = pytest.mark.gpu


def orc_ssmh(w, X, y, seed, step0, n_steps, block_mode, Nbatch, k0, eps, delta_mode=0, delta_eps=0.0):
    w = np.ascontiguousarray(w.copy())
    D, P = w.shape
    stats = np.empty((P, 3))
    oracle.lib().orc_blr_subsampled_mh(oracle._ptr(w), oracle._ptr(X), oracle._ptr(y), oracle.I32(D),
                                       oracle.I64(X.shape[0]), oracle.I64(P), oracle.I64(0), oracle.F64(0.0),
                                       oracle.F64(1.0), oracle.U64(seed), oracle.U32(step0), oracle.I32(n_steps),
                                       oracle.I32(block_mode), oracle.F64(Nbatch), oracle.I32(k0), oracle.F64(eps),
                                       oracle.I32(delta_mode), oracle.F64(delta_eps), oracle._ptr(stats))
    return w, stats


@pytest.mark.parametrize("N,D,P,Nbatch,k0,eps,block,mode", [
    (1000, 5, 40, 64, 3, 0.05, "one", -1),
    (777, 16, 64, 50, 2, 0.01, "one", -1),
    (300, 3, 33, 7, 3, 0.1, 1, 1),
    (500, 4, 20, 100, 1, 0.0, "one", -1),      # epsilon = 0: runs to n == N, exact MH decision
])
def test_subsampled_mh_matches_oracle(N, D, P, Nbatch, k0, eps, block, mode):
    X, y = make_data(N, D, seed=6)
    ts = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), P, seed=61)
    ts.observe(X, y)
    ts.incorporate()
    w0 = read_w(ts, D)
    K = 25
    step0 = ts.rng_step
    stats = ts.primitive_infer(["subsampled_mh", "default", block, Nbatch, k0, eps, False, 0, False, K])
    w_ref, stats_ref = orc_ssmh(w0, X, y, 61, step0, K, mode, Nbatch, k0, eps)
    np.testing.assert_allclose(read_w(ts, D), w_ref, rtol=1e-12)
    np.testing.assert_array_equal(stats[:, 0], stats_ref[:, 0])          # accepted transitions
    np.testing.assert_allclose(stats[:, 1], stats_ref[:, 1], rtol=1e-12)  # mean sections visited
    fin = np.isfinite(stats_ref[:, 2])
    np.testing.assert_array_equal(np.isfinite(stats[:, 2]), fin)
    np.testing.assert_allclose(stats[fin, 2], stats_ref[fin, 2], rtol=1e-8)
    assert stats[:, 0].sum() > 0
    if eps > 0:
        assert stats[:, 1].mean() < N       # early stopping actually happens
    # the stale logit cache is refreshed before the next exact operation
    np.testing.assert_allclose(ts.log_likelihood_at("default", "all"), orc_loglik(w_ref, X, y), rtol=1e-12)


def test_drift_kernel_extension_matches_oracle():
    N, D, P = 400, 4, 32
    X, y = make_data(N, D, seed=7)
    ts = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), P, seed=62)
    ts.observe(X, y)
    ts.incorporate()
    w0 = read_w(ts, D)
    step0 = ts.rng_step
    from venturecxx_b200 import _lib
    st = np.empty((P, 3))
    _lib.call("vk_subsampled_mh", ts._h, 0, _lib.BLOCK_ONE, 50, 3, 0.05, 2, 0.3, 0, 40, None, _lib.ptr(st))
    w_ref, st_ref = orc_ssmh(w0, X, y, 62, step0, 40, -1, 50, 3, 0.05, delta_mode=2, delta_eps=0.3)
    np.testing.assert_allclose(read_w(ts, D), w_ref, rtol=1e-12)
    np.testing.assert_array_equal(st[:, 0], st_ref[:, 0])


def test_feistel_is_a_permutation():
    N = 1000
    out = np.empty(N, dtype=np.uint64)
    oracle.lib().orc_feistel_perm(oracle.U64(5), oracle.U32(3), oracle.U64(N), oracle._ptr(out))
    assert sorted(out.tolist()) == list(range(N))


def test_use_delta_flag_is_inert_for_normal_sites():
    # NormalOutputPSP.hasDeltaKernel() is False (continuous.py:352-353), so loadKernels
    # (scaffold.py:298-310) installs no delta kernel: useDeltaKernels=true changes nothing
    N, D, P = 200, 3, 16
    X, y = make_data(N, D, seed=9)
    res = []
    for flag in (False, True):
        ts = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), P, seed=63)
        ts.observe(X, y)
        ts.incorporate()
        ts.primitive_infer(["subsampled_mh", "default", "one", 20, 3, 0.05, flag, 0.5, False, 10])
        res.append(read_w(ts, D))
    np.testing.assert_array_equal(res[0], res[1])
