"""GPU parity of the fused SP primitives, through the batched PSP surface,
against the oracle on the same Philox streams, the reference-generated golden
values, and the KS / chi-square checks of test/conformance/sps/."""
import numpy as np
import pytest
import scipy.stats as st

import oracle
from helpers import SP, close, golden, unj
from venturecxx_b200.sp import BatchArgs, builtin_psp

pytestmark = pytest.mark.gpu
PS = golden("lite_psps.json")
U = golden("lite_utils.json")


def test_log_density_golden_values():
    n = builtin_psp("normal")
    c = PS["normal"]
    got = n.logDensity([k["x"] for k in c], BatchArgs([[k["mu"] for k in c], [k["sigma"] for k in c]]))
    assert all(close(g, k["out"], rtol=1e-13, atol=1e-14) for g, k in zip(got, c))
    c = PS["gamma"]
    got = builtin_psp("gamma").logDensity([k["x"] for k in c], BatchArgs([[k["shape"] for k in c], [k["rate"] for k in c]]))
    assert all(close(g, unj(k["out"]), rtol=1e-12, atol=1e-13) for g, k in zip(got, c))
    c = PS["beta"]
    got = builtin_psp("beta").logDensity([k["x"] for k in c], BatchArgs([[k["a"] for k in c], [k["b"] for k in c]]))
    assert all(close(g, unj(k["out"]), rtol=1e-12, atol=1e-13) for g, k in zip(got, c))
    for key, name, pk in [("flip", "flip", "p"), ("log_flip", "log_flip", "logp"), ("log_odds_flip", "log_odds_flip", "x")]:
        c = PS[key]
        got = builtin_psp(name).logDensity([k["val"] for k in c], BatchArgs([[k[pk] for k in c]]))
        assert all(close(g, unj(k["out"]), rtol=1e-13) for g, k in zip(got, c)), key
    for c in U["logDensityCategorical"]:
        got = builtin_psp("categorical").logDensity([c["val"]], BatchArgs([[c["ps"]]]))[0]
        assert close(got, unj(c["out"]), rtol=1e-13)
    for c in U["logDensityDirichlet"]:
        got = builtin_psp("dirichlet").logDensity([c["theta"]], BatchArgs([[c["alpha"]]]))[0]
        assert close(got, c["out"], rtol=1e-12)
    for c in PS["sym_dir_cat"]:
        psp = builtin_psp("make_sym_dir_cat")
        got = psp.logDensity(np.arange(c["n"]), BatchArgs([[c["alpha"]]], aux=np.tile(c["counts"], (c["n"], 1))))
        np.testing.assert_allclose(got, c["logDensity"], rtol=1e-13)
    for c in PS["beta_bernoulli"]:
        psp = builtin_psp("make_beta_bernoulli")
        got = psp.logDensity([1, 0], BatchArgs([[c["a"]], [c["b"]]], aux=[[c["yes"], c["no"]]] * 2))
        np.testing.assert_allclose(got, [c["logDensity_true"], c["logDensity_false"]], rtol=1e-13)


@pytest.mark.parametrize("name,ops", [
    ("normal", [1.5, 2.0]), ("gamma", [10.0, 10.0]), ("gamma", [0.3, 2.0]), ("beta", [2.0, 3.0]),
    ("beta", [0.5, 0.5]), ("beta", [0.2, 1.5]), ("flip", [0.3]), ("log_flip", [-0.7]), ("log_odds_flip", [0.4])])
def test_simulate_matches_oracle_stream(name, ops):
    n = 20000
    args = BatchArgs([np.full(n, o) for o in ops], seed=77, step=5)
    got = builtin_psp(name).simulate(args)
    ref = oracle.sp_simulate(SP[name], np.tile(ops, (n, 1)), n, seed=77, step=5)
    np.testing.assert_allclose(got, ref, rtol=1e-12, atol=1e-14)
    dens = builtin_psp(name).logDensity(got, args)
    np.testing.assert_allclose(dens, oracle.sp_log_density(SP[name], ref, np.tile(ops, (n, 1)), n), rtol=1e-10, atol=1e-12)


def test_vector_and_count_sps_match_oracle():
    n = 5000
    alpha = [0.4, 1.0, 2.5, 0.7]
    got = builtin_psp("dirichlet").simulate(BatchArgs([np.tile(alpha, (n, 1))], seed=3, step=1))
    ref = oracle.sp_simulate(SP["dirichlet"], np.tile(alpha, (n, 1)), n, dim=4, seed=3, step=1)
    np.testing.assert_allclose(got, ref, rtol=1e-12)
    np.testing.assert_allclose(got.sum(1), 1.0, rtol=1e-12)
    got = builtin_psp("symmetric_dirichlet").simulate(BatchArgs([np.full(n, 0.5), [6]], seed=3, step=2))
    ref = oracle.sp_simulate(SP["symmetric_dirichlet"], np.full((n, 1), 0.5), n, dim=6, seed=3, step=2)
    np.testing.assert_allclose(got, ref, rtol=1e-12)
    ps = [0.1, 0.0, 0.5, 0.4]
    got = builtin_psp("categorical").simulate(BatchArgs([np.tile(ps, (n, 1))], seed=4))
    assert np.array_equal(got, oracle.sp_simulate(SP["categorical"], np.tile(ps, (n, 1)), n, seed=4))
    counts = np.tile([3, 0, 1, 5, 0], (n, 1))
    got = builtin_psp("make_sym_dir_cat").simulate(BatchArgs([[0.5]], seed=5, aux=counts))
    ref = oracle.sp_simulate(SP["sym_dir_cat"], np.concatenate([np.full((n, 1), 0.5), counts], 1), n, dim=5, seed=5)
    assert np.array_equal(got, ref)
    # range_tree.sample frequencies from the reference (dirichlet.py:164-166)
    r = PS["range_tree_sample"]
    mine = np.bincount(got.astype(int), minlength=5)
    from stats_policy import report_same_discrete_counts, report_test
    report_test(report_same_discrete_counts(mine + 1, np.array(r["freq"]) + 1))


def test_sampler_distributions():
    from stats_policy import report_known_continuous, statistical_test
    n = 100000

    @statistical_test
    def gamma_ks(seed):                                              # test_gamma.py:26-34
        g = builtin_psp("gamma").simulate(BatchArgs([np.full(n, 10.0), np.full(n, 10.0)], seed=seed))
        return report_known_continuous(st.gamma(10, scale=0.1).cdf, g, "gamma(10, 10)")
    gamma_ks()
    for a, b in [(2.0, 3.0), (0.5, 0.5), (0.3, 0.3), (1.0, 1.0)]:
        @statistical_test
        def beta_ks(seed):                                           # test_beta.py:147-170
            x = builtin_psp("beta").simulate(BatchArgs([np.full(n, a), np.full(n, b)], seed=seed))
            assert np.all((x >= 0) & (x <= 1))
            return report_known_continuous(st.beta(a, b).cdf, x, "beta(%g, %g)" % (a, b))
        beta_ks()
    f = builtin_psp("flip").simulate(BatchArgs([np.full(n, 0.3)], seed=13))
    assert abs(f.mean() - 0.3) < 5 * np.sqrt(0.21 / n)
    # degenerate spikes, continuous.py:737-742
    assert np.all(builtin_psp("beta").simulate(BatchArgs([np.zeros(8), np.ones(8)], seed=1)) == 1.0)
    assert np.all(builtin_psp("beta").simulate(BatchArgs([np.ones(8), np.zeros(8)], seed=1)) == 0.0)
