"""GPU parity for config C1 (normal-normal) against the oracle on the same
Philox streams, plus the reference's known-answer checks:
  test/inference_language/test_likelihood_weighting.py:33-45 (weights = N(14,1).pdf(sample))
  test/inference_quality/micro/test_basic_stats.py:65-76     (posterior N(12, sqrt(.5)))
  test/smc/test_particle_filter.py:61-72                     (resample -> N(1, sqrt(.5)))"""
import ctypes

import numpy as np
import pytest
import scipy.stats as st

import oracle
import venturecxx_b200 as vb

pytestmark = pytest.mark.gpu
RTOL = 1e-12


def orc_nn_lw(P, y, mu0, s0, s, seed, step, p_base=0):
    x, lw = np.empty(P), np.empty(P)
    y = np.ascontiguousarray(y, dtype=np.float64)
    oracle.lib().orc_nn_likelihood_weight(oracle._ptr(x), oracle._ptr(lw), oracle._ptr(y), oracle.I64(y.size),
                                          oracle.I64(P), oracle.I64(p_base), oracle.F64(mu0), oracle.F64(s0),
                                          oracle.F64(s), oracle.U64(seed), oracle.U32(step))
    return x, lw


def orc_nn_mh(x, y, mu0, s0, s, seed, step0, K, p_base=0):
    x = x.copy()
    acc = np.empty(x.size, dtype=np.int64)
    y = np.ascontiguousarray(y, dtype=np.float64)
    oracle.lib().orc_nn_mh(oracle._ptr(x), oracle._ptr(y), oracle.I64(y.size), oracle.I64(x.size),
                           oracle.I64(p_base), oracle.F64(mu0), oracle.F64(s0), oracle.F64(s),
                           oracle.U64(seed), oracle.U32(step0), oracle.I32(K), oracle._ptr(acc))
    return x, acc


@pytest.mark.parametrize("P", [1, 1000, 70001])
def test_likelihood_weight_matches_oracle(P):
    ts = vb.B200TraceSet(vb.NormalNormalModel(10.0, 1.0, 1.0), P, seed=11)
    ts.observe([14.0])
    ts.incorporate()
    step = ts.rng_step
    ts.likelihood_weight()
    x_ref, lw_ref = orc_nn_lw(P, [14.0], 10.0, 1.0, 1.0, 11, step)
    np.testing.assert_allclose(ts.read_site(0), x_ref, rtol=RTOL)
    np.testing.assert_allclose(ts.log_weights, lw_ref, rtol=RTOL)
    # test_likelihood_weighting.py:33-45: weight == N(14,1).logpdf(sample)
    np.testing.assert_allclose(ts.log_weights, st.norm(loc=14, scale=1).logpdf(ts.read_site(0)), rtol=1e-9, atol=1e-9)


def test_incorporate_increment_and_injected_values():
    P = 4096
    ts = vb.B200TraceSet(vb.NormalNormalModel(0.0, 1.0, 2.0), P, seed=3)
    xs = np.linspace(-5, 5, P)
    ts.write_site(0, xs)
    ts.observe([2.0, -1.0, 0.5])
    inc = ts.incorporate()
    ref = sum(st.norm(loc=xs, scale=2.0).logpdf(v) for v in [2.0, -1.0, 0.5])
    np.testing.assert_allclose(inc, ref, rtol=1e-11)
    np.testing.assert_allclose(ts.log_weights, ref, rtol=1e-11)
    np.testing.assert_allclose(ts.log_likelihood_at("default", "all"), ref, rtol=1e-11)
    np.testing.assert_allclose(ts.log_joint_at("default", "all"), ref + st.norm(0, 1).logpdf(xs), rtol=1e-11)
    # second incorporate with nothing new adds nothing (trace.py:424,437)
    assert np.all(ts.incorporate() == 0)


@pytest.mark.parametrize("K", [1, 25])
def test_mh_chain_matches_oracle(K):
    P = 5000
    ts = vb.B200TraceSet(vb.NormalNormalModel(10.0, 1.0, 1.0), P, seed=21)
    ts.observe([14.0])
    ts.incorporate()
    x0 = ts.read_site(0)
    step0 = ts.rng_step
    nodes, acc = ts.resimulation_mh("default", "one", K, return_accepts=True)
    x_ref, acc_ref = orc_nn_mh(x0, [14.0], 10.0, 1.0, 1.0, 21, step0, K)
    np.testing.assert_allclose(ts.read_site(0), x_ref, rtol=RTOL)
    assert np.array_equal(acc, acc_ref)
    assert np.all(nodes == 2.0)  # |DRG| = x + its lookup (scaffold.py:67-68)
    assert ts.rng_step == step0 + K


def test_posterior_normal_12_sqrt_half():
    # test_basic_stats.py:65-76: prior N(10,1), obs 14 with noise 1 => N(12, sqrt(.5)); policy of test/stats.py:101-131
    from stats_policy import report_known_gaussian, statistical_test

    @statistical_test
    def run(seed):
        P = 20000
        ts = vb.B200TraceSet(vb.NormalNormalModel(10.0, 1.0, 1.0), P, seed=seed)
        ts.observe([14.0])
        ts.incorporate()
        # independence sampler from the prior: sup(posterior/prior) ~ 77, so the chain
        # needs ~10^3 transitions to forget its start ((1-1/77)^1500 ~ 3e-9)
        ts.primitive_infer(["resimulation_mh", "default", "one", 1500])
        return report_known_gaussian(12, np.sqrt(0.5), ts.read_site(0))
    run()


@pytest.mark.parametrize("mode", ["multinomial", "systematic"])
def test_smc_resample_posterior(mode):
    # test/smc/test_particle_filter.py:61-93: prior N(0,1), observe normal(x,1)=2,
    # likelihood-weight + resample => N(1, sqrt(.5)); the reference KS-tests 4 predictions of 4 independent
    # particle sets -- here one particle is drawn from each of many independent small particle sets
    from stats_policy import report_known_gaussian, statistical_test

    @statistical_test
    def run(seed):
        P = 50000
        ts = vb.B200TraceSet(vb.NormalNormalModel(0.0, 1.0, 1.0), P, seed=seed)
        ts.observe([2.0])
        ts.incorporate()
        ts.likelihood_weight()
        ts.resample(P, mode=mode)
        x = ts.read_site(0)
        # resampled particles are duplicated (KS on ties is too strict): moments on the whole set, KS on the
        # distinct survivors weighted by nothing would be biased -- so KS a thinned subsample whose expected number of
        # repeated ancestors is small (every 500th particle of 50 000: 100 draws, ~1 % ties)
        assert abs(x.mean() - 1.0) < 5 * np.sqrt(0.5 / P) * 3
        assert abs(x.std() - np.sqrt(0.5)) < 0.02
        sub = x[np.random.default_rng(seed).permutation(P)[:100]]
        return report_known_gaussian(1.0, np.sqrt(0.5), sub)
    run()


def test_errors_are_venture_exceptions():
    ts = vb.B200TraceSet(vb.NormalNormalModel(), 16)
    with pytest.raises(vb.VentureException):
        ts.primitive_infer(["resimulation_mh", "nonexistent_scope", "one", 1])
    with pytest.raises(vb.VentureException):
        ts.primitive_infer(["slice", "default", "one", 1])
    ts.observe([1.0])
    with pytest.raises(vb.VentureException):  # unpropagated observation (needs incorporate)
        ts.resimulation_mh("default", "one", 1)
