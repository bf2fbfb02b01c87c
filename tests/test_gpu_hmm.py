"""GPU parity for config C3 (examples/hmm.vnt particle filter) against the
oracle and the reference's filtered marginals (examples/hmm_plugin.py:175-205)."""
import numpy as np
import pytest

import oracle
import venturecxx_b200 as vb
from helpers import OracleHMM, golden

pytestmark = pytest.mark.gpu
H = golden("hmm.json")


def make(P, seed, T_max=20):
    return vb.B200TraceSet(vb.HMMModel(H["transition_matrix"], H["emission_matrix"], 1, T_max), P, seed=seed)


def hist_of(ts, T):
    return np.stack([ts.read_site(t) for t in range(T)], axis=1).astype(np.uint8)


@pytest.mark.parametrize("mode", ["systematic", "multinomial"])
def test_filter_matches_oracle_step_by_step(mode):
    P = 3001
    ts = make(P, 31)
    o = OracleHMM(H["transition_matrix"], H["emission_matrix"], 1, P, 32, seed=31)
    ts.rng_step = 0
    for t, d in enumerate(H["data"][:8]):
        ts.observe(d)
        s = ts.rng_step
        inc = ts.incorporate()
        np.testing.assert_allclose(inc, o.step(d, s), rtol=1e-12)
        np.testing.assert_allclose(ts.log_weights, o.lw, rtol=1e-12)
        assert np.array_equal(hist_of(ts, t + 1), o.hist[:, :t + 1])
        s = ts.rng_step
        anc = ts.resample(P, mode=mode, return_ancestors=True)
        ref_anc = oracle.ancestors(o.lw, P, 1 if mode == "systematic" else 0, seed=31, step=s)
        assert np.array_equal(anc, ref_anc)
        o.resample(ref_anc)
        assert np.array_equal(hist_of(ts, t + 1), o.hist[:, :t + 1])   # whole mem table moved with the particle
        np.testing.assert_allclose(ts.log_weights, o.lw, rtol=1e-12)


def test_filtered_marginals_match_reference_table():
    # particle_filter(P) of hmm.vnt:77-92; metric3 of hmm_plugin.py:130-132,225-232
    P = 1 << 18
    ts = make(P, 5)
    tvd = []
    for t, d in enumerate(H["data"]):
        ts.observe(d)
        ts.incorporate()
        ts.resample(P, mode="systematic")
        m = np.bincount(ts.read_site(t).astype(int), minlength=6)[1:6] / P
        tvd.append(0.5 * np.abs(m - np.array(H["filtered_marginals"][t])).sum())
    assert np.mean(tvd) < 0.004, (np.mean(tvd), tvd)


def test_likelihood_weight_and_mh_match_oracle():
    P = 2000
    ts = make(P, 41)
    o = OracleHMM(H["transition_matrix"], H["emission_matrix"], 1, P, 32, seed=41)
    for d in H["data"][:10]:
        ts.observe(d)
        o.step(d, ts.rng_step)
        ts.incorporate()
    s = ts.rng_step
    ts.likelihood_weight()
    o.prior(s)
    assert np.array_equal(hist_of(ts, 10), o.hist[:, :10])
    np.testing.assert_allclose(ts.log_weights, o.lw, rtol=1e-12)
    # the emission of the deterministic state(1) is not on the border of scaffold(default, all)
    # (Lite: tests/golden/lite_chains.json c3 "ll_at"), every other emission is
    e0 = np.log(np.array(H["emission_matrix"])[0, H["data"][0] - 1])
    np.testing.assert_allclose(ts.log_likelihood_at("default", "all"), o.lw - e0, rtol=1e-12)
    assert ts.numBlocksInScope() == 9
    for block, mode, n_steps in [("one", -1, 30), ("each", -3, 2 * 9), (3, 3, 5)]:
        s = ts.rng_step
        K = 30 if block == "one" else (2 if block == "each" else 5)
        _, acc = ts.resimulation_mh("default", block, K, return_accepts=True)
        ref_acc = o.mh(s, n_steps, mode)
        assert np.array_equal(hist_of(ts, 10), o.hist[:, :10])
        assert np.array_equal(acc, ref_acc)
