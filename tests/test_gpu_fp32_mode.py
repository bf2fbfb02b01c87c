"""fp32 throughput mode of the logistic-regression MH loop (SURVEY 8 preamble: "fp64 (parity) and fp32
(throughput) modes"; north_star: log densities within 1e-5 relative in fp32 mode).

The reference is double throughout (utils.py:257-268 log_logistic on Python floats), so fp32 mode is judged
against the fp64 path and the oracle: (1) the running log likelihood the fp32 kernel leaves behind is within
1e-5 relative of the fp64 value at the same weights; (2) on the same Philox streams the fp32 chain makes the
same accept decisions as the oracle's chain except at near-ties; (3) the posterior it samples has the moments
the fp64 path samples; (4) the logit cache is recomputed in fp64 before any fp64 consumer reads it."""
import numpy as np
import pytest

import venturecxx_b200 as vb
from test_gpu_blr import make_data, orc_loglik, orc_mh, read_w

pytestmark = pytest.mark.gpu
FP32_RTOL = 1e-5      # north_star's tolerance for fp32 mode


def raw_buffer(ts, which, n):
    """Device doubles as they are, without the refresh an API read would trigger."""
    from cuda.bindings import runtime as cudart
    ptr, nbytes = ts.device_buffer(which)
    assert nbytes >= 8 * n
    out = np.empty(n, dtype=np.float64)
    ts.synchronize()
    err, = cudart.cudaMemcpy(out.ctypes.data, ptr, 8 * n, cudart.cudaMemcpyKind.cudaMemcpyDeviceToHost)
    assert int(err) == 0
    return out


def blr(N, D, P, seed, precision, data_seed=4):
    X, y = make_data(N, D, seed=data_seed)
    ts = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), P, seed=seed, precision=precision)
    ts.observe(X, y)
    ts.incorporate()
    return ts, X, y


def test_precision_switch():
    ts, X, y = blr(50, 3, 16, 3, "fp64")
    assert ts.precision == "fp64"
    ts.precision = "fp32"
    assert ts.precision == "fp32"
    with pytest.raises(vb.VentureException):
        ts.precision = "fp16"
    for model in [vb.NormalNormalModel(), vb.LDAModel(2, 3, 1, 4)]:    # no fp32 kernel: refused, not ignored
        with pytest.raises(vb.VentureException):
            vb.B200TraceSet(model, 8, seed=1, precision="fp32")


@pytest.mark.parametrize("N,D,P,K,block", [(1000, 8, 2048, 64, "one"), (4097, 5, 512, 6, "each"), (333, 4, 700, 40, 2),
                                           (10000, 16, 256, 160, "one")])
def test_fp32_log_density_within_tolerance(N, D, P, K, block):
    ts, X, y = blr(N, D, P, 31, "fp32")
    nodes, acc = ts.resimulation_mh("default", block, K, return_accepts=True)
    assert acc.sum() > 0
    ll32 = raw_buffer(ts, 3, P)                       # rho as the fp32 kernel left it
    w = read_w(ts, D)
    ll64 = orc_loglik(w, X, y)
    np.testing.assert_allclose(ll32, ll64, rtol=FP32_RTOL)
    assert np.max(np.abs(ll32 - ll64) / np.abs(ll64)) < 2e-6     # what the kernel achieves; tolerance has slack
    # in fp32 mode every log density is promised to 1e-5; back in fp64 the cache is rebuilt from the fp64 weights
    np.testing.assert_allclose(ts.log_likelihood_at("default", "all"), ll64, rtol=FP32_RTOL)
    ts.precision = "fp64"
    np.testing.assert_allclose(ts.log_likelihood_at("default", "all"), ll64, rtol=1e-12)
    per = D if block == "each" else 1
    assert np.all(nodes == (3 + 2 * N) * per)


def test_fp32_chain_follows_the_oracle_chain():
    # same (seed, particle, step) streams => same proposals; accept decisions can differ only where
    # log u is within the fp32 error of xi - rho
    N, D, P, K = 1000, 8, 4096, 64
    ts, X, y = blr(N, D, P, 37, "fp32")
    w0 = read_w(ts, D)
    step0 = ts.rng_step
    nodes, acc = ts.resimulation_mh("default", "one", K, return_accepts=True)
    w_ref, acc_ref, ll_ref = orc_mh(w0, X, y, 37, step0, K, -1)
    w = read_w(ts, D)
    same = np.all(np.isclose(w, w_ref, rtol=1e-12, atol=0.0), axis=0)     # proposals are fp64 draws
    assert same.mean() > 0.995, same.mean()
    assert np.array_equal(acc[same], acc_ref[same])
    assert ts.rng_step == step0 + K
    # several launches in a row stay on the chain too (the float row is carried between launches)
    ts2, _, _ = blr(N, D, P, 37, "fp32")
    for _ in range(4):
        ts2.resimulation_mh("default", "one", K // 4)
    assert np.all(np.isclose(read_w(ts2, D), w_ref, rtol=1e-12, atol=0.0), axis=0).mean() > 0.995


def test_fp32_posterior_moments_match_fp64():
    N, D, P = 200, 3, 1 << 15
    out = {}
    for prec in ("fp64", "fp32"):
        ts, X, y = blr(N, D, P, 41, prec, data_seed=9)
        ts.resimulation_mh("default", "each", 150)
        out[prec] = read_w(ts, D)
    a, b = out["fp64"], out["fp32"]
    se = a.std(axis=1) / np.sqrt(P / 20.0)            # prior proposals at N = 200: integrated autocorrelation ~20
    assert np.all(np.abs(a.mean(axis=1) - b.mean(axis=1)) < 4 * se)
    np.testing.assert_allclose(a.std(axis=1), b.std(axis=1), rtol=0.05)


def test_fp32_log_weights_within_tolerance():
    """north_star: log densities AND log weights within 1e-5 in fp32 mode -- likelihood_weight, incorporate
    (makeConsistent) and log_likelihood_at of the logistic-regression model run their single-precision variant"""
    N, D, P = 10000, 16, 512
    X, y = make_data(N, D, seed=3)
    ts = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), P, seed=53, precision="fp32")
    w = read_w(ts, D)
    ts.observe(X[:6000], y[:6000])
    inc1 = ts.incorporate()
    np.testing.assert_allclose(inc1, orc_loglik(w, X, y, 0, 6000), rtol=FP32_RTOL)
    ts.observe(X[6000:], y[6000:])
    inc2 = ts.incorporate()
    np.testing.assert_allclose(inc2, orc_loglik(w, X, y, 6000, N), rtol=FP32_RTOL)
    np.testing.assert_allclose(ts.log_weights, orc_loglik(w, X, y), rtol=FP32_RTOL)
    assert np.max(np.abs(ts.log_weights - orc_loglik(w, X, y)) / np.abs(orc_loglik(w, X, y))) < 2e-6
    ts.likelihood_weight()
    w2 = read_w(ts, D)
    assert not np.array_equal(w, w2)
    np.testing.assert_allclose(ts.log_weights, orc_loglik(w2, X, y), rtol=FP32_RTOL)
    np.testing.assert_allclose(ts.log_likelihood_at("default", "all"), orc_loglik(w2, X, y), rtol=FP32_RTOL)
    ts.precision = "fp64"                               # the same handle back in parity mode
    np.testing.assert_allclose(ts.log_likelihood_at("default", "all"), orc_loglik(w2, X, y), rtol=1e-12)
    ts.likelihood_weight()
    np.testing.assert_allclose(ts.log_weights, orc_loglik(read_w(ts, D), X, y), rtol=1e-12)


def test_fp32_then_resample_and_incorporate_stay_exact():
    # after fp32 calls: resample carries the state, incorporate is the fp32 variant; back in fp64 everything is exact
    N, D, P = 600, 4, 256
    X, y = make_data(N, D, seed=6)
    ts = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), P, seed=43, precision="fp32")
    ts.observe(X[:400], y[:400])
    ts.incorporate()
    ts.resimulation_mh("default", "one", 30)
    anc = ts.resample(P, mode="systematic", return_ancestors=True)
    w = read_w(ts, D)
    ts.observe(X[400:], y[400:])
    incr = ts.incorporate()
    np.testing.assert_allclose(incr, orc_loglik(w, X, y, 400, N), rtol=FP32_RTOL)
    np.testing.assert_allclose(ts.log_likelihood_at("default", "all"), orc_loglik(w, X, y), rtol=FP32_RTOL)
    ts.precision = "fp64"                              # and the fp64 kernel picks the chain up exactly
    np.testing.assert_allclose(ts.log_likelihood_at("default", "all"), orc_loglik(w, X, y), rtol=1e-12)
    step0 = ts.rng_step
    ts.resimulation_mh("default", "one", 8)
    w_ref, _, ll_ref = orc_mh(w, X, y, 43, step0, 8, -1)
    np.testing.assert_allclose(read_w(ts, D), w_ref, rtol=1e-12)
    np.testing.assert_allclose(ts.log_likelihood_at("default", "all"), ll_ref, rtol=1e-12)


def test_fp32_rows_too_long_for_shared_memory_run_fp64():
    N, D, P, K = 60000, 3, 8, 4
    ts, X, y = blr(N, D, P, 47, "fp32", data_seed=8)
    w0 = read_w(ts, D)
    step0 = ts.rng_step
    ts.resimulation_mh("default", "one", K)
    w_ref, _, ll_ref = orc_mh(w0, X, y, 47, step0, K, -1)
    np.testing.assert_allclose(read_w(ts, D), w_ref, rtol=1e-12)
    np.testing.assert_allclose(ts.log_likelihood_at("default", "all"), ll_ref, rtol=1e-12)
