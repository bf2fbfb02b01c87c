"""CPU: the oracle (oracle/oracle.c) against fixtures produced by executing the
reference's own functions (tests/golden/make_golden.py) and against the literal
tables in the reference's tests."""
import math

import numpy as np
import pytest
import scipy.stats as st

import oracle
from helpers import SP, OracleHMM, close, golden, hmm_forward_marginals, unj

U = golden("lite_utils.json")
PS = golden("lite_psps.json")


def test_logsumexp_golden():
    for c in U["logsumexp"]:
        got = oracle.logsumexp(np.array(unj(c["in"]), dtype=np.float64))
        assert close(got, unj(c["out"]), rtol=1e-14), c


def test_logistic_family_golden():
    L = oracle.lib()
    for c in U["log_logistic"]:
        assert close(L.orc_log_logistic(unj(c["x"])), unj(c["out"]), rtol=1e-15), c
    for c in U["logistic"]:
        assert close(L.orc_logistic(unj(c["x"])), unj(c["out"]), rtol=1e-15), c
    # test/unit/test_logistic.py:66-130 table
    for x, Lx, logLx in [(0, 0.5, math.log(0.5)), (-1, 0.2689414213699951, -1.3132616875182228),
                         (1, 0.7310585786300049, -0.3132616875182228), (710, 1, -4.4762862256751298e-309),
                         (-710, 4.4762862256751298e-309, -710), (1000, 1, 0), (-1000, 0, -1000), (1e308, 1, 0),
                         (-1e308, 0, -1e308)]:
        assert close(L.orc_logistic(float(x)), Lx, rtol=1e-15)
        assert close(L.orc_log_logistic(float(x)), logLx, rtol=1e-15)


def test_weights_golden():
    for c in U["log_domain_even_out"]:
        a = np.array(unj(c["in"]))
        got = oracle.logsumexp(a) - math.log(c["n"])
        assert all(close(got, v, rtol=1e-14) for v in unj(c["out"]))
    for c in U["logWeightsToNormalizedDirect"]:
        got = oracle.normalized_probs(np.array(unj(c["in"])))
        assert all(close(g, v, rtol=1e-14) for g, v in zip(got, unj(c["out"])))


def test_categorical_golden():
    for c in U["logDensityCategorical"]:
        got = oracle.sp_log_density(SP["categorical"], [c["val"]], [c["ps"]], 1)[0]
        assert close(got, unj(c["out"]), rtol=1e-14), c
    for c in U["logDensityLogCategorical"]:
        lps = [math.log(p) if p > 0 else -math.inf for p in c["ps"]]
        got = oracle.sp_log_density(SP["log_categorical"], [c["val"]], [lps], 1)[0]
        assert close(got, unj(c["out"]), rtol=1e-13, atol=1e-15), c
    # repeated objects: the reference sums the mass of equal objects (utils.py:112-118); the compiled
    # models only use distinct objects, for which this is the plain categorical
    c = U["logDensityCategorical_objects"]
    assert close(unj(c["out"])[0], math.log(0.1 + 0.3), rtol=1e-14)


def test_dirichlet_golden():
    for c in U["logDensityDirichlet"]:
        n = len(c["alpha"])
        got = oracle.sp_log_density(SP["dirichlet"], c["theta"], [c["alpha"]], 1, dim=n)[0]
        assert close(got, c["out"], rtol=1e-12), c


def test_normal_gamma_beta_golden():
    for c in PS["normal"]:
        got = oracle.sp_log_density(SP["normal"], [c["x"]], [[c["mu"], c["sigma"]]], 1)[0]
        assert close(got, c["out"], rtol=1e-14, atol=1e-14)
    for c in PS["gamma"]:
        got = oracle.sp_log_density(SP["gamma"], [c["x"]], [[c["shape"], c["rate"]]], 1)[0]
        assert close(got, unj(c["out"]), rtol=1e-12, atol=1e-13), c
    for c in PS["beta"]:
        got = oracle.sp_log_density(SP["beta"], [c["x"]], [[c["a"], c["b"]]], 1)[0]
        assert close(got, unj(c["out"]), rtol=1e-12, atol=1e-13), c


def test_flip_family_golden():
    for key, spn, pk in [("flip", "flip", "p"), ("log_flip", "log_flip", "logp"), ("log_odds_flip", "log_odds_flip", "x")]:
        for c in PS[key]:
            got = oracle.sp_log_density(SP[spn], [c["val"]], [[c[pk]]], 1)[0]
            assert close(got, unj(c["out"]), rtol=1e-14), (key, c)


def test_dir_cat_and_beta_bernoulli_golden():
    L = oracle.lib()
    for c in PS["sym_dir_cat"]:
        n = c["n"]
        params = [[c["alpha"]] + [float(v) for v in c["counts"]]]
        for v in range(n):
            got = oracle.sp_log_density(SP["sym_dir_cat"], [v], params, 1, dim=n)[0]
            assert close(got, c["logDensity"][v], rtol=1e-13)
        counts = np.array(c["counts"], dtype=np.float64)
        got = L.orc_sym_dircat_log_density_of_data(oracle._ptr(counts), oracle.I32(n), oracle.F64(c["alpha"]))
        assert close(got, c["logDensityOfData"], rtol=1e-12, atol=1e-13)
    for c in PS["beta_bernoulli"]:
        q = [[c["a"], c["b"], float(c["yes"]), float(c["no"])]]
        assert close(oracle.sp_log_density(SP["beta_bernoulli"], [1], q, 1)[0], c["logDensity_true"], rtol=1e-13)
        assert close(oracle.sp_log_density(SP["beta_bernoulli"], [0], q, 1)[0], c["logDensity_false"], rtol=1e-13)
        got = L.orc_beta_bernoulli_log_density_of_data(c["a"], c["b"], float(c["yes"]), float(c["no"]))
        assert close(got, c["logDensityOfData"], rtol=1e-12, atol=1e-13)
    # test/conformance/test_get_global_logscore.py:32-50
    g = PS["global_logscore"]
    assert close(100 * oracle.sp_log_density(SP["flip"], [1], [[0.5]], 1)[0], g["flip_half_x100"], rtol=1e-14)
    tot = sum(oracle.sp_log_density(SP["beta_bernoulli"], [1], [[1.0, 1.0, float(k), 0.0]], 1)[0] for k in range(100))
    assert close(tot, g["beta_bernoulli_1_1_100_true"], rtol=1e-12)
    assert close(L.orc_beta_bernoulli_log_density_of_data(1.0, 1.0, 100.0, 0.0), g["beta_bernoulli_1_1_100_true"], rtol=1e-12)


def test_sequential_test_golden():
    G = golden("sequential_test.json")
    L = oracle.lib()
    for c in G["t_cdf_scipy"]:
        assert close(L.orc_student_t_cdf(c["t"], float(c["df"])), c["out"], rtol=1e-10, atol=1e-15), c
    import ctypes
    for c in G["cases"]:
        d = np.array(c["dllh"], dtype=np.float64)
        n, t = ctypes.c_double(), ctypes.c_double()
        acc = L.orc_sequential_test(oracle.F64(c["mu_0"]), oracle.I32(c["k0"]), oracle.F64(c["Nbatch"]), oracle.F64(c["N"]),
                                    oracle.F64(c["epsilon"]), oracle._ptr(d), ctypes.byref(n), ctypes.byref(t))
        assert acc == c["accept"] and n.value == c["n"], c
        assert close(t.value, unj(c["tstat"]), rtol=1e-9), c
        if "tstat_expect_reference_test" in c:  # test_subsampledmh.py:56-60
            assert abs(t.value - c["tstat_expect_reference_test"]) < 1e-5


def test_samplers_are_distributionally_right():
    n = 40000
    assert st.kstest(oracle.sp_simulate(SP["normal"], np.tile([1.5, 2.0], (n, 1)), n, seed=1), st.norm(1.5, 2.0).cdf).pvalue > 1e-4
    # test/conformance/sps/test_gamma.py:26-34: gamma(10,10)
    assert st.kstest(oracle.sp_simulate(SP["gamma"], np.tile([10.0, 10.0], (n, 1)), n, seed=2), st.gamma(10, scale=0.1).cdf).pvalue > 1e-4
    assert st.kstest(oracle.sp_simulate(SP["gamma"], np.tile([0.3, 2.0], (n, 1)), n, seed=3), st.gamma(0.3, scale=0.5).cdf).pvalue > 1e-4
    for a, b in [(2.0, 3.0), (0.5, 0.5), (0.2, 1.5)]:
        assert st.kstest(oracle.sp_simulate(SP["beta"], np.tile([a, b], (n, 1)), n, seed=4), st.beta(a, b).cdf).pvalue > 1e-4
    ps = [0.1, 0.0, 0.5, 0.4]
    draws = oracle.sp_simulate(SP["categorical"], np.tile(ps, (n, 1)), n, seed=5).astype(int)
    obs = np.bincount(draws, minlength=4)
    assert obs[1] == 0
    assert st.chisquare(obs[[0, 2, 3]], n * np.array([0.1, 0.5, 0.4])).pvalue > 1e-4


def test_resampler_matches_reference_multinomial_frequencies():
    # the reference resamples with sampleLogCategorical (trace_set.py:209-219); 20000 of its draws are in the fixture
    c = U["sampleLogCategorical"]
    lw = np.array(unj(c["logs"]))
    anc = oracle.ancestors(lw, c["n"], 0, seed=9, step=0)
    mine = np.bincount(anc, minlength=len(lw))
    ref = np.array(c["counts"])
    assert mine[3] == 0 and ref[3] == 0
    keep = [0, 1, 2, 4]
    assert st.chi2_contingency(np.array([mine[keep], ref[keep]]))[1] > 1e-4
    sys_anc = oracle.ancestors(lw, c["n"], 1, seed=9, step=0)
    p = np.exp(lw - lw.max()); p /= p.sum()
    assert np.all(np.abs(np.bincount(sys_anc, minlength=5) - c["n"] * p) <= 1.0 + 1e-9)  # systematic: within one of n*p


def test_hmm_fixture_and_oracle_filter():
    H = golden("hmm.json")
    exact = hmm_forward_marginals(H["transition_matrix"], H["emission_matrix"], H["data"], H["init_state"])
    np.testing.assert_allclose(exact, np.array(H["filtered_marginals"]), atol=6e-9)  # hmm_plugin.py:175-205 (8 decimals)
    P = 20000
    o = OracleHMM(H["transition_matrix"], H["emission_matrix"], 1, P, 32, seed=3)
    tvd = []
    for t, d in enumerate(H["data"]):
        o.step(d, rng_step=2 * t)
        anc = oracle.ancestors(o.lw, P, 1, seed=3, step=2 * t + 1)
        o.resample(anc)
        m = np.bincount(o.hist[:, t], minlength=6)[1:6] / P
        tvd.append(0.5 * np.abs(m - exact[t]).sum())
    assert np.mean(tvd) < 0.01, np.mean(tvd)
