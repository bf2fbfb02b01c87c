"""GPU, at BASELINE.json's full sizes: size-independent properties of the hot path
(the oracle cannot finish these sizes in seconds, so parity is checked through
invariants the domain offers: cache idempotence, systematic-resampling counts,
count-table consistency, a posterior computed by quadrature)."""
import json
import os

import numpy as np
import pytest

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


def test_c2_bench_shape_matches_oracle_on_sampled_particles():
    """The launch bench.py times -- P = 65 536, N = 10^4, D = 16, (resimulation_mh default one 160), the
    two-proposal shared-memory kernel -- against orc_blr_mh on sampled particles (the oracle's p_base picks
    the particle's Philox stream): weights, accept counts and log likelihoods at 1e-12."""
    import oracle
    N, D, P, T = 10000, 16, 65536, 160
    X, y, _ = blr_data(N, D)
    ts = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), P, seed=1234)
    ts.observe(X, y)
    ts.incorporate()
    sample = [0, 1, 255, 256, 4095, 30001, 50000, 65534, 65535]
    w0 = np.stack([ts.read_site(j) for j in range(D)])
    step0 = ts.rng_step
    _, acc = ts.resimulation_mh("default", "one", T, return_accepts=True)
    w1 = np.stack([ts.read_site(j) for j in range(D)])
    ll1 = ts.log_likelihood_at("default", "all")
    for p in sample:
        w = np.ascontiguousarray(w0[:, p:p + 1].copy())
        a, ll = np.empty(1, dtype=np.int64), np.empty(1)
        oracle.lib().orc_blr_mh(oracle._ptr(w), oracle._ptr(X), oracle._ptr(y), oracle.I32(D), oracle.I64(N), oracle.I64(1),
                                oracle.I64(p), oracle.F64(0.0), oracle.F64(1.0), oracle.U64(1234), oracle.U32(step0),
                                oracle.I32(T), oracle.I32(-1), oracle._ptr(a), oracle._ptr(ll))
        np.testing.assert_allclose(w1[:, p], w[:, 0], rtol=1e-12, err_msg="particle %d" % p)
        assert acc[p] == a[0], p
        np.testing.assert_allclose(ll1[p], ll[0], rtol=1e-12)
    assert acc[sample].sum() > 0


def test_long_rows_flush_the_running_products():
    """N = 70 000 > 64 tiles of 1024: the `(it & 63) == 63` flush of the running products (a product of 64
    factors in [1,2] stays below 2^64) runs in the streaming kernel; chain equal to the oracle's."""
    from test_gpu_blr import make_data, orc_mh, read_w
    N, D, P, K = 70000, 3, 6, 5
    X, y = make_data(N, D, seed=18)
    ts = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), P, seed=29)
    ts.observe(X, y)
    ts.incorporate()
    w0 = read_w(ts, D)
    step0 = ts.rng_step
    _, acc = ts.resimulation_mh("default", "one", K, return_accepts=True)
    w_ref, acc_ref, ll_ref = orc_mh(w0, X, y, 29, step0, K, -1)
    np.testing.assert_allclose(read_w(ts, D), w_ref, rtol=1e-12)
    assert np.array_equal(acc, acc_ref)
    np.testing.assert_allclose(ts.log_likelihood_at("default", "all"), ll_ref, rtol=1e-12)
    # fp32 mode asks for a float row in shared memory; a row this long runs the fp64 kernel instead (same chain)
    ts.precision = "fp32"
    step1 = ts.rng_step
    ts.resimulation_mh("default", "one", K)
    w_ref2, _, _ = orc_mh(w_ref, X, y, 29, step1, K, -1)
    np.testing.assert_allclose(read_w(ts, D), w_ref2, rtol=1e-12)


def _d1_posterior_cdf(X, y):
    grid = np.linspace(-6, 6, 24001)
    eta = grid[:, None] * X[:, 0][None, :]
    s = np.where(y[None, :] == 1, 1.0, -1.0)
    logp = -0.5 * grid ** 2 + (-np.log1p(np.exp(-s * eta))).sum(axis=1)
    p = np.exp(logp - logp.max())
    cdf = np.cumsum(p); cdf /= cdf[-1]
    return lambda v: np.interp(v, grid, cdf)


def test_subsampled_mh_posterior_matches_quadrature_d1():
    """The reference pins subsampled_mh statistically (test/inference_language/test_subsampledmh.py:64-97: KS of
    the chain's samples against the exact posterior, N = 100 observations, Nbatch = 10, k0 = 3, eps = 0.01).
    Same here on the D = 1 logistic-regression posterior computed by quadrature.  This also tests the
    documented deviation that all particles of a transition share one minibatch order."""
    from stats_policy import report_known_continuous, statistical_test
    N, D, P = 100, 1, 4000
    X, y, _ = blr_data(N, D, seeds=(20, 21, 22))
    cdf = _d1_posterior_cdf(X, y)

    @statistical_test
    def run(seed):
        ts = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), P, seed=seed)
        ts.observe(X, y)
        ts.incorporate()
        st_ = ts.subsampled_mh("default", "one", 10, 3, 0.01, False, 0, False, 200)
        assert st_[:, 1].mean() < N                     # the sequential test does stop early on average
        return report_known_continuous(cdf, ts.read_site(0)[::2], "the quadrature posterior")
    run()


def test_blr_posterior_matches_quadrature_d1():
    """Statistical parity under the reference's policy (test/stats.py:101-131, tests/stats_policy.py): D = 1
    logistic regression, posterior of w computed by quadrature on a fine grid."""
    from stats_policy import report_known_continuous, statistical_test
    N, D, P = 40, 1, 20000
    X, y, _ = blr_data(N, D, seeds=(10, 11, 12))
    cdf = _d1_posterior_cdf(X, y)

    @statistical_test
    def run(seed):
        ts = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), P, seed=seed)
        ts.observe(X, y)
        ts.incorporate()
        ts.resimulation_mh("default", "one", 300)
        return report_known_continuous(cdf, ts.read_site(0), "the quadrature posterior")
    run()


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
