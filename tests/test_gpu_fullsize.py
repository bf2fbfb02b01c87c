"""GPU, at BASELINE.json's full sizes: size-independent properties of the hot path
(the oracle cannot finish these sizes in seconds, so parity is checked through
invariants the domain offers: cache idempotence, systematic-resampling counts,
count-table consistency, a posterior computed by quadrature)."""
import json
import os

import numpy as np
import pytest
import scipy.stats as st

import venturecxx_b200 as vb
from helpers import golden

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def blr_data(N, D, seeds=(0, 1, 2)):
    X = np.random.default_rng(seeds[0]).normal(size=(N, D))
    w_true = np.random.default_rng(seeds[1]).normal(size=D)
    y = (np.random.default_rng(seeds[2]).random(N) < 1.0 / (1.0 + np.exp(-X @ w_true))).astype(np.uint8)
    return X, y, w_true


def test_c2_full_size_cache_is_idempotent_and_chain_moves_to_the_data():
    """C2: P = 65 536, N = 10^4, D = 16, (resimulation_mh default one 160).  After the launch the
    incrementally maintained logit cache / log-likelihood must equal a from-scratch recomputation
    from the weights (refresh), to 1e-10 relative; accept counts are consistent with the chain."""
    N, D, P, T = 10000, 16, 65536, 160
    X, y, w_true = blr_data(N, D)
    ts = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), P, seed=1234)
    ts.observe(X, y)
    inc = ts.incorporate()
    ll0 = ts.log_likelihood_at("default", "all")
    np.testing.assert_allclose(inc, ll0, rtol=1e-12)
    nodes, acc = ts.resimulation_mh("default", "one", T, return_accepts=True)
    assert np.all(nodes == 3 + 2 * N)                    # |DRG| of scaffold(w_j), dispatch.py:109-120
    assert acc.min() >= 0 and acc.max() <= T and 0.005 < acc.mean() / T < 0.5   # ~0.2 while burning in from the prior, ~0.03 at stationarity
    ll_cached = ts.log_likelihood_at("default", "all")
    ts.refresh()                                          # recompute eta and ll from w
    ll_fresh = ts.log_likelihood_at("default", "all")
    np.testing.assert_allclose(ll_cached, ll_fresh, rtol=1e-10)
    # prior draws fit the data terribly; every accepted move must have been uphill or nearly so
    assert np.median(ll_fresh) > np.median(ll0)
    moved = acc > 0
    assert np.all(ll_fresh[~moved] == ll0[~moved])        # rejected-only particles are untouched
    # a second launch continues the same chains deterministically
    s = ts.rng_step
    _, acc2 = ts.resimulation_mh("default", "one", 40, return_accepts=True)
    w_a = ts.read_site(3).copy()
    ts2 = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), 4096, seed=1234)
    ts2.observe(X, y); ts2.incorporate()
    ts2.resimulation_mh("default", "one", T)
    ts2.rng_step = s
    ts2.resimulation_mh("default", "one", 40)
    np.testing.assert_array_equal(ts2.read_site(3), w_a[:4096])   # streams are keyed by particle, not by batch shape


def test_blr_posterior_matches_quadrature_d1():
    """Statistical parity (test/stats.py policy: KS at p > 0.001 with one retry): D = 1 logistic
    regression, posterior of w computed by quadrature on a fine grid."""
    N, D, P = 40, 1, 20000
    X, y, _ = blr_data(N, D, seeds=(10, 11, 12))
    grid = np.linspace(-6, 6, 24001)
    eta = grid[:, None] * X[:, 0][None, :]
    s = np.where(y[None, :] == 1, 1.0, -1.0)
    logp = -0.5 * grid ** 2 + (-np.log1p(np.exp(-s * eta))).sum(axis=1)
    p = np.exp(logp - logp.max())
    cdf = np.cumsum(p); cdf /= cdf[-1]
    for attempt in range(2):
        ts = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), P, seed=77 + attempt)
        ts.observe(X, y)
        ts.incorporate()
        ts.resimulation_mh("default", "one", 300)
        w = ts.read_site(0)
        pv = st.kstest(w, lambda v: np.interp(v, grid, cdf)).pvalue
        if pv > 0.001:
            break
    assert pv > 0.001, pv


def test_c3_full_size_systematic_resampling_properties():
    """C3: P = 2^20.  Systematic resampling: ancestors non-decreasing, offspring count of particle i
    within 1 of P*w_i, new weights = logsumexp - log P; all -inf weights => identity-like uniform."""
    P = 1 << 20
    H = golden("hmm.json")
    ts = vb.B200TraceSet(vb.HMMModel(H["transition_matrix"], H["emission_matrix"], 1, 20), P, seed=3)
    for d in H["data"][:3]:
        ts.observe(d)
    ts.incorporate()
    lw = ts.log_weights
    lse = ts.logsumexp()
    m = lw.max()
    assert abs(lse - (m + np.log(np.exp(lw - m).sum()))) < 1e-10
    anc = ts.resample(P, mode="systematic", return_ancestors=True)
    assert np.all(np.diff(anc) >= 0) and anc.min() >= 0 and anc.max() < P
    counts = np.bincount(anc, minlength=P)
    expect = P * np.exp(lw - lse)
    assert counts.sum() == P and np.all(np.abs(counts - expect) < 1.0 + 1e-6)
    np.testing.assert_allclose(ts.log_weights, lse - np.log(P), rtol=1e-12)
    # multinomial mode: counts are Binomial(P, w_i); check the first two moments loosely over states
    ts.set_particle_log_weights(lw)
    anc = ts.resample(P, mode="multinomial", return_ancestors=True)
    counts = np.bincount(anc, minlength=P)
    assert counts.sum() == P
    z = (counts - expect) / np.sqrt(np.maximum(expect, 1e-12))
    assert abs(z.mean()) < 0.01 and 0.9 < z.var() < 1.1
    # all particles impossible: uniform resampling (utils.py:95-99,176-180)
    ts.set_particle_log_weights(np.full(P, -np.inf))
    anc = ts.resample(P, mode="systematic", return_ancestors=True)
    assert np.array_equal(anc, np.arange(P))
    assert np.all(np.isneginf(ts.log_weights))


def test_c4_full_size_count_tables_stay_consistent():
    """C4: K = 16, V = 1024, 64 docs x 64 tokens, P = 4096.  After MH sweeps the incrementally
    maintained count tables must equal a recount from z (refresh): the AAA weights of both alpha
    sites, which read every table, are unchanged by the recount."""
    from test_gpu_lda import corpus
    K, V, n_docs, n_per, P = 16, 1024, 64, 64, 4096
    doc, word = corpus(K, V, n_docs, n_per)
    n_tok = len(doc)
    ts = vb.B200TraceSet(vb.LDAModel(K, V, n_docs, n_tok), P, seed=7)
    ts.observe(doc, word)
    ts.incorporate()
    _, acc = ts.resimulation_mh("default", "one", 2 * (n_tok + 2), return_accepts=True)
    assert 0.05 < acc.mean() / (2 * (n_tok + 2)) <= 1.0
    a0, a1 = ts.log_likelihood_at("default", 0), ts.log_likelihood_at("default", 1)
    z5 = ts.read_site(2 + 5).copy()
    ts.refresh()
    np.testing.assert_allclose(ts.log_likelihood_at("default", 0), a0, rtol=1e-12)
    np.testing.assert_allclose(ts.log_likelihood_at("default", 1), a1, rtol=1e-12)
    assert np.array_equal(ts.read_site(2 + 5), z5) and z5.min() >= 0 and z5.max() < K
    # resampling moves whole rows: the same invariants hold afterwards
    ts.set_particle_log_weights(np.random.default_rng(0).normal(size=P))
    anc = ts.resample(P, mode="systematic", return_ancestors=True)
    np.testing.assert_allclose(ts.log_likelihood_at("default", 1), a1[anc], rtol=1e-12)


def test_c5_subsampled_mh_visits_few_sections_and_matches_exact_mh_decisions():
    """C5 shape at N = 10^6 on one GPU: with epsilon = 0.01 most proposals are decided after a few
    minibatches; with epsilon = 0 the test runs to n == N and is the exact MH decision."""
    N, D, P = 10 ** 6, 16, 64
    X, y, _ = blr_data(N, D, seeds=(4, 5, 6))
    ts = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), P, seed=9)
    ts.observe(X, y)
    ts.incorporate()
    stats = ts.subsampled_mh("default", "one", 65536, 3, 0.01, False, 0, False, 10)
    assert stats[:, 1].mean() < 0.6 * N and np.all(stats[:, 1] >= 3 * 65536 - 1e-9)
    # epsilon = 0: every test runs to n == N (the exact MH decision, subsampled_mh.py:150)
    ts_a = vb.B200TraceSet(vb.LogisticRegressionModel(D, 20000), P, seed=9)
    ts_a.observe(X[:20000], y[:20000]); ts_a.incorporate()
    sa = ts_a.subsampled_mh("default", "one", 4096, 1, 0.0, False, 0, False, 25)
    assert np.all(sa[:, 1] == 20000)
    w_a = np.stack([ts_a.read_site(j) for j in range(D)])
    assert np.all(np.isfinite(w_a)) and sa[:, 0].sum() > 0
