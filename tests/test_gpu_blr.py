"""GPU parity for config C2 (Bayesian logistic regression, single-site MH)
against the oracle, which recomputes every vector_dot and every log_odds_flip
logDensity from scratch as the reference's regen does
(regen.py:120-127,255-285; discrete.py:148-150; utils.py:257-268)."""
import numpy as np
import pytest

import oracle
import venturecxx_b200 as vb

pytestmark = pytest.mark.gpu


def make_data(N, D, seed=0):
    X = np.random.default_rng(seed).normal(size=(N, D))
    w_true = np.random.default_rng(seed + 1).normal(size=D)
    p = 1 / (1 + np.exp(-X @ w_true))
    y = (np.random.default_rng(seed + 2).random(N) < p).astype(np.uint8)
    return X, y


def orc_loglik(w, X, y, i0=0, i1=None):
    D, P = w.shape
    i1 = X.shape[0] if i1 is None else i1
    out = np.empty(P)
    oracle.lib().orc_blr_log_likelihood(oracle._ptr(w), oracle._ptr(X), oracle._ptr(y), oracle.I32(D),
                                        oracle.I64(i0), oracle.I64(i1), oracle.I64(P), oracle._ptr(out))
    return out


def orc_prior(D, P, seed, step, mu0=0.0, s0=1.0):
    w = np.empty((D, P))
    oracle.lib().orc_blr_prior(oracle._ptr(w), oracle.I32(D), oracle.I64(P), oracle.I64(0), oracle.F64(mu0),
                               oracle.F64(s0), oracle.U64(seed), oracle.U32(step))
    return w


def orc_mh(w, X, y, seed, step0, n_steps, block_mode, mu0=0.0, s0=1.0):
    w = np.ascontiguousarray(w.copy())
    D, P = w.shape
    acc, ll = np.empty(P, dtype=np.int64), np.empty(P)
    oracle.lib().orc_blr_mh(oracle._ptr(w), oracle._ptr(X), oracle._ptr(y), oracle.I32(D),
                            oracle.I64(X.shape[0]), oracle.I64(P), oracle.I64(0), oracle.F64(mu0),
                            oracle.F64(s0), oracle.U64(seed), oracle.U32(step0), oracle.I32(n_steps),
                            oracle.I32(block_mode), oracle._ptr(acc), oracle._ptr(ll))
    return w, acc, ll


def read_w(ts, D):
    return np.stack([ts.read_site(j) for j in range(D)])


@pytest.mark.parametrize("N,D,P", [(1, 1, 3), (257, 5, 300), (1000, 16, 64)])
def test_likelihood_weight_and_incorporate(N, D, P):
    X, y = make_data(N, D)
    ts = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), P, seed=13)
    w0 = read_w(ts, D)
    np.testing.assert_allclose(w0, orc_prior(D, P, 13, 0), rtol=1e-12)
    # observe in two batches; incorporate adds each batch's log-likelihood
    h = N // 2
    ts.observe(X[:h], y[:h])
    inc1 = ts.incorporate()
    ts.observe(X[h:], y[h:])
    inc2 = ts.incorporate()
    np.testing.assert_allclose(inc1, orc_loglik(w0, X, y, 0, h), rtol=1e-12, atol=1e-13)
    np.testing.assert_allclose(inc2, orc_loglik(w0, X, y, h, N), rtol=1e-12, atol=1e-13)
    np.testing.assert_allclose(ts.log_weights, orc_loglik(w0, X, y), rtol=1e-12)
    step = ts.rng_step
    ts.likelihood_weight()
    w1 = read_w(ts, D)
    np.testing.assert_allclose(w1, orc_prior(D, P, 13, step), rtol=1e-12)
    np.testing.assert_allclose(ts.log_weights, orc_loglik(w1, X, y), rtol=1e-12)
    np.testing.assert_allclose(ts.log_likelihood_at("default", "all"), orc_loglik(w1, X, y), rtol=1e-12)


@pytest.mark.parametrize("block,mode,mult", [("one", -1, 1), ("each", -3, None), (2, 2, 1), ("all", -2, 1)])
def test_mh_chain_matches_oracle(block, mode, mult):
    N, D, P, K = 400, 6, 200, 12
    X, y = make_data(N, D, seed=4)
    ts = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), P, seed=17)
    ts.observe(X, y)
    ts.incorporate()
    w0 = read_w(ts, D)
    step0 = ts.rng_step
    nodes, acc = ts.resimulation_mh("default", block, K, return_accepts=True)
    n_steps = K * D if block == "each" else K
    w_ref, acc_ref, ll_ref = orc_mh(w0, X, y, 17, step0, n_steps, mode)
    np.testing.assert_allclose(read_w(ts, D), w_ref, rtol=1e-12)
    assert np.array_equal(acc, acc_ref)
    assert acc.sum() > 0
    np.testing.assert_allclose(ts.log_likelihood_at("default", "all"), ll_ref, rtol=1e-12)
    assert ts.rng_step == step0 + n_steps


def test_mh_global_memory_path_matches_shared_memory_path():
    # rows too large for shared memory take the streaming kernel; same chain
    N, D, P, K = 30000, 3, 8, 6
    X, y = make_data(N, D, seed=8)
    ts = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), P, seed=19)
    ts.observe(X, y)
    ts.incorporate()
    w0 = read_w(ts, D)
    step0 = ts.rng_step
    ts.resimulation_mh("default", "one", K)
    w_ref, _, ll_ref = orc_mh(w0, X, y, 19, step0, K, -1)
    np.testing.assert_allclose(read_w(ts, D), w_ref, rtol=1e-12)
    np.testing.assert_allclose(ts.log_likelihood_at("default", "all"), ll_ref, rtol=1e-12)


def test_resample_carries_logit_cache():
    N, D, P = 300, 4, 128
    X, y = make_data(N, D, seed=2)
    ts = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), P, seed=23)
    ts.observe(X, y)
    ts.incorporate()
    w0 = read_w(ts, D)
    anc = ts.resample(P, mode="systematic", return_ancestors=True)
    np.testing.assert_array_equal(read_w(ts, D), w0[:, anc])
    # the cached logits moved with their particle: an MH step afterwards agrees with the oracle
    step0 = ts.rng_step
    ts.resimulation_mh("default", "one", 5)
    w_ref, _, ll_ref = orc_mh(w0[:, anc], X, y, 23, step0, 5, -1)
    np.testing.assert_allclose(read_w(ts, D), w_ref, rtol=1e-12)
    np.testing.assert_allclose(ts.log_likelihood_at("default", "all"), ll_ref, rtol=1e-12)
