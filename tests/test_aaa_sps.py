"""A10 / A12: the remaining Dirichlet / Bernoulli makers -- make_dir_cat (asymmetric, collapsed), make_uc_dir_cat,
make_uc_sym_dir_cat, make_uc_beta_bernoulli -- their made SPs and AAA LKernels, against
tests/golden/lite_aaa.json, produced by calling the reference's own classes (tests/golden/make_aaa_golden.py).

CPU: the oracle (oracle/oracle.c) against the golden.  GPU: the batched SP entry points (vk_sp_*) against the
golden and the oracle; draws on the oracle's Philox streams must come out as the same parameters.
"""
import numpy as np
import pytest

import oracle
from helpers import golden

G = golden("lite_aaa.json")
RTOL = 1e-12
DIR_CAT, UC_DIR_CAT, SUFF, SYM_DIR_CAT, BETA_BERN, DIRICHLET, SYM_DIRICHLET, BETA = 13, 14, 15, 11, 12, 9, 10, 3


def rows(case):
    n = case["n"]
    return np.array([case["alpha"] + case["counts"]], dtype=np.float64), n


def test_oracle_matches_reference_classes():
    seed, step = G["seed"], G["step"]
    for c in G["dir_cat"]:
        p, n = rows(c)
        pn = np.tile(p, (n, 1))
        np.testing.assert_allclose(oracle.sp_log_density(DIR_CAT, np.arange(n, dtype=np.float64), pn, n, dim=n), c["collapsed_logDensity"], rtol=RTOL)
        np.testing.assert_allclose(oracle.sp_log_density_of_data(DIR_CAT, p, 1, dim=n), c["collapsed_logDensityOfData"], rtol=RTOL)
        np.testing.assert_allclose(c["marginalLogDensityOfData"], c["collapsed_logDensityOfData"], rtol=RTOL)
        batch = np.tile(np.array(c["alpha"]), (c["row"] + 1, 1))
        theta = oracle.sp_simulate(DIRICHLET, batch, c["row"] + 1, dim=n, seed=seed, step=step)[c["row"]]
        np.testing.assert_allclose(theta, c["maker_theta"], rtol=RTOL)
        np.testing.assert_allclose(oracle.sp_log_density(DIRICHLET, np.array([theta]), np.array([c["alpha"]]), 1, dim=n), c["maker_logDensity"], rtol=RTOL)
        tn = np.tile(np.array(c["maker_theta"]), (n, 1))
        np.testing.assert_allclose(oracle.sp_log_density(UC_DIR_CAT, np.arange(n, dtype=np.float64), tn, n, dim=n), c["uncollapsed_logDensity"], rtol=RTOL)
        th, w = oracle.sp_aaa_simulate(UC_DIR_CAT, np.tile(p, (c["row"] + 1, 1)), c["row"] + 1, dim=n, seed=seed, step=step + 1)
        np.testing.assert_allclose(th[c["row"]], c["aaa_theta"], rtol=RTOL)
        np.testing.assert_allclose(w[c["row"]], c["aaa_weight"], rtol=RTOL)
    for c in G["sym_dir_cat"]:
        n = c["n"]
        p = np.array([[c["alpha"]] + c["counts"]], dtype=np.float64)
        np.testing.assert_allclose(oracle.sp_log_density_of_data(SYM_DIR_CAT, p, 1, dim=n), c["marginalLogDensityOfData"], rtol=RTOL)
        th, w = oracle.sp_aaa_simulate(SYM_DIRICHLET, np.tile(p, (c["row"] + 1, 1)), c["row"] + 1, dim=n, seed=seed, step=step + 1)
        np.testing.assert_allclose(th[c["row"]], c["aaa_theta"], rtol=RTOL)
        np.testing.assert_allclose(w[c["row"]], c["aaa_weight"], rtol=RTOL)
        theta = oracle.sp_simulate(SYM_DIRICHLET, np.full((c["row"] + 1, 1), c["alpha"]), c["row"] + 1, dim=n, seed=seed, step=step)[c["row"]]
        np.testing.assert_allclose(theta, c["maker_theta"], rtol=RTOL)
        np.testing.assert_allclose(oracle.sp_log_density(SYM_DIRICHLET, np.array([theta]), np.array([[c["alpha"]]]), 1, dim=n), c["maker_logDensity"], rtol=RTOL)
    for c in G["beta_bernoulli"]:
        p = np.array([[c["a"], c["b"], c["yes"], c["no"]]], dtype=np.float64)
        np.testing.assert_allclose(oracle.sp_log_density_of_data(BETA_BERN, p, 1), c["marginalLogDensityOfData"], rtol=RTOL, atol=1e-15)
        wt, w = oracle.sp_aaa_simulate(SUFF, np.tile(p, (c["row"] + 1, 1)), c["row"] + 1, seed=seed, step=step + 1, vector=False)
        np.testing.assert_allclose(wt[c["row"]], c["aaa_weight_param"], rtol=RTOL)
        np.testing.assert_allclose(w[c["row"]], c["aaa_weight"], rtol=RTOL, atol=1e-15)
        q = np.array([[c["aaa_weight_param"], c["yes"], c["no"]]])
        np.testing.assert_allclose(oracle.sp_log_density_of_data(SUFF, q, 1), c["suff_logDensityOfData"], rtol=RTOL, atol=1e-15)
        np.testing.assert_allclose(oracle.sp_log_density(SUFF, np.array([1.0, 0.0]), np.tile(q[:, :1], (2, 1)), 2),
                                   [c["suff_logDensity_true"], c["suff_logDensity_false"]], rtol=RTOL)
        mw = oracle.sp_simulate(BETA, np.tile([c["a"], c["b"]], (c["row"] + 1, 1)), c["row"] + 1, seed=seed, step=step)[c["row"]]
        np.testing.assert_allclose(mw, c["maker_weight"], rtol=RTOL)
        np.testing.assert_allclose(oracle.sp_log_density(BETA, np.array([mw]), np.array([[c["a"], c["b"]]]), 1), c["maker_logDensity"], rtol=1e-10)


@pytest.mark.gpu
def test_device_sps_match_reference_classes():
    from venturecxx_b200.sp import AAALKernel, BatchArgs, builtin_psp
    seed, step = G["seed"], G["step"]
    for c in G["dir_cat"]:
        n, r = c["n"], c["row"]
        alpha, counts = np.array(c["alpha"]), np.array(c["counts"], dtype=np.float64)
        psp = builtin_psp("make_dir_cat")
        args = BatchArgs([alpha], aux=np.tile(counts, (n, 1)))
        np.testing.assert_allclose(psp.logDensity(np.arange(n), args), c["collapsed_logDensity"], rtol=RTOL)
        np.testing.assert_allclose(psp.logDensityOfData(BatchArgs([alpha], aux=counts)), c["collapsed_logDensityOfData"], rtol=RTOL)
        upsp = builtin_psp("make_uc_dir_cat")
        np.testing.assert_allclose(upsp.logDensity(np.arange(n), BatchArgs([np.tile(c["maker_theta"], (n, 1))])), c["uncollapsed_logDensity"], rtol=RTOL)
        theta = builtin_psp("dirichlet").simulate(BatchArgs([np.tile(alpha, (r + 1, 1))], seed=seed, step=step))[r]
        np.testing.assert_allclose(theta, c["maker_theta"], rtol=RTOL)
        k = AAALKernel("make_uc_dir_cat")
        th, w = k.simulate(BatchArgs([alpha], seed=seed, step=step + 1, aux=np.tile(counts, (r + 1, 1))), return_weight=True)
        np.testing.assert_allclose(th[r], c["aaa_theta"], rtol=RTOL)
        np.testing.assert_allclose(w[r], c["aaa_weight"], rtol=RTOL)
        np.testing.assert_allclose(k.weight(BatchArgs([alpha], aux=counts)), c["aaa_weight"], rtol=RTOL)
        # frequencies of the made SPs' simulate (dirichlet.py:164-166, 271-273) by chi-square under the reference's policy
        from stats_policy import report_known_counts, report_test
        m = 40000
        draws = psp.simulate(BatchArgs([alpha], seed=9, aux=np.tile(counts, (m, 1))))
        report_test(report_known_counts(alpha + counts, np.bincount(draws.astype(int), minlength=n)))
        draws = upsp.simulate(BatchArgs([np.tile(c["maker_theta"], (m, 1))], seed=10))
        report_test(report_known_counts(c["maker_theta"], np.bincount(draws.astype(int), minlength=n)))
    for c in G["sym_dir_cat"]:
        n, r = c["n"], c["row"]
        counts = np.array(c["counts"], dtype=np.float64)
        k = AAALKernel("make_uc_sym_dir_cat")
        th, w = k.simulate(BatchArgs([np.array([c["alpha"]])], seed=seed, step=step + 1, aux=np.tile(counts, (r + 1, 1))), return_weight=True)
        np.testing.assert_allclose(th[r], c["aaa_theta"], rtol=RTOL)
        np.testing.assert_allclose(w[r], c["aaa_weight"], rtol=RTOL)
    for c in G["beta_bernoulli"]:
        r = c["row"]
        k = AAALKernel("make_uc_beta_bernoulli")
        aux = np.tile([c["yes"], c["no"]], (r + 1, 1)).astype(np.float64)
        wt, w = k.simulate(BatchArgs([np.array([c["a"]]), np.array([c["b"]])], seed=seed, step=step + 1, aux=aux), return_weight=True)
        np.testing.assert_allclose(wt[r], c["aaa_weight_param"], rtol=RTOL)
        np.testing.assert_allclose(w[r], c["aaa_weight"], rtol=RTOL, atol=1e-15)
        spsp = builtin_psp("make_uc_beta_bernoulli")
        a1 = BatchArgs([np.array([c["aaa_weight_param"]])])
        np.testing.assert_allclose(spsp.logDensity(np.array([1.0]), a1), c["suff_logDensity_true"], rtol=RTOL)
        np.testing.assert_allclose(spsp.logDensity(np.array([0.0]), a1), c["suff_logDensity_false"], rtol=RTOL)
        np.testing.assert_allclose(spsp.logDensityOfData(BatchArgs([np.array([c["aaa_weight_param"]])], aux=[[c["yes"], c["no"]]])),
                                   c["suff_logDensityOfData"], rtol=RTOL, atol=1e-15)
