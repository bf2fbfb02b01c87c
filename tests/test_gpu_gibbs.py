"""Enumerative Gibbs, `(gibbs scope block transitions)`: the CUDA path through the C ABI
against the oracle on the same Philox streams (the oracle itself is pinned against Lite's
EnumerativeGibbsOperator in tests/test_oracle_lite_chains.py), against the Lite-run fixture directly, and
against exact posteriors: the HMM's smoothed marginals (examples/hmm_plugin.py:130-174, metric of
:225-232) and a brute-force enumeration of a tiny LDA."""
import itertools
import math

import numpy as np
import pytest

import venturecxx_b200 as vb
from helpers import OracleHMM, OracleLDA, golden, unj
from test_gpu_lda import corpus, state_of

pytestmark = pytest.mark.gpu
H = golden("hmm.json")
CH = golden("lite_chains.json")


def hmm_ts(P, seed, T_max=32):
    return vb.B200TraceSet(vb.HMMModel(H["transition_matrix"], H["emission_matrix"], 1, T_max), P, seed=seed)


def hist_of(ts, T):
    return np.stack([ts.read_site(t) for t in range(T)], axis=1).astype(np.uint8)


def test_hmm_gibbs_matches_oracle():
    P, T = 3000, 12
    ts = hmm_ts(P, 71)
    o = OracleHMM(H["transition_matrix"], H["emission_matrix"], 1, P, 32, seed=71)
    for d in H["data"][:T]:
        ts.observe(d)
        o.step(d, ts.rng_step)
        ts.incorporate()
    for block, mode, K, n_steps in [("one", -1, 40, 40), ("each", -3, 2, 2 * (T - 1)), (4, 4, 6, 6), (T - 2, T - 2, 3, 3)]:
        s = ts.rng_step
        nodes, ch = ts.gibbs("default", block, K, return_changes=True)
        ref = o.gibbs(s, n_steps, mode)
        assert np.array_equal(hist_of(ts, T), o.hist[:, :T]), block
        assert np.array_equal(ch, ref)
        assert ts.rng_step == s + n_steps
    assert np.all(nodes == 9.0)       # the newest state's scaffold (lite_chains.json c3)
    # zero transitions: nothing moves, no step consumed
    s = ts.rng_step
    ts.primitive_infer(["gibbs", "default", "one", 0])
    assert ts.rng_step == s


def test_hmm_gibbs_vs_lite_fixture():
    c = CH["c3"]
    recs, T, K = c["particles"], c["T"], c["K"]
    ts = vb.B200TraceSet(vb.HMMModel(c["trans"], c["emit"], 1, 32), c["P"], seed=c["seed"])
    for t in range(T):
        ts.observe(c["data"][t])
        ts.rng_step = c["steps"]["filter"] + t
        ts.incorporate()
    for t in range(1, T):          # Lite's state when its Gibbs chain began
        ts.write_site(t, np.array([r["hist_each"][t] for r in recs], dtype=np.float64))
    ts.rng_step = c["steps"]["gibbs_one"]
    nodes = ts.primitive_infer(["gibbs", "default", "one", K])
    assert np.array_equal(hist_of(ts, T), [r["hist_gibbs_one"] for r in recs])
    np.testing.assert_allclose(nodes, [np.mean(r["nodes_gibbs_one"]) for r in recs], rtol=1e-15)
    ts.rng_step = c["steps"]["gibbs_each"]
    ts.primitive_infer(["gibbs", "default", "each", 2])
    assert np.array_equal(hist_of(ts, T), [r["hist_gibbs_each"] for r in recs])


def test_hmm_gibbs_sweeps_reach_smoothed_marginals():
    # hmm.vnt observations, then ordered Gibbs sweeps over state(2..20): the particles' marginals must
    # approach the exact smoothed marginals tabulated in examples/hmm_plugin.py (TVD, :225-232)
    P, T = 1 << 16, 20
    ts = hmm_ts(P, 73)
    for d in H["data"]:
        ts.observe(d)
        ts.incorporate(return_increments=False)
    ts.primitive_infer(["gibbs", "default", "each", 3000])   # single-site Gibbs on a chain mixes slowly
    hist = hist_of(ts, T)
    sm = np.array(H["smoothed_marginals"])
    tvd = []
    for t in range(T):
        emp = np.bincount(hist[:, t], minlength=6)[1:6] / P
        tvd.append(0.5 * np.abs(emp - sm[t]).sum())
    assert max(tvd) < 0.02, tvd


@pytest.mark.parametrize("K,V,n_docs,n_per,P", [(2, 2, 3, 4, 200), (16, 40, 4, 16, 150), (40, 7, 2, 9, 64)])
def test_lda_gibbs_matches_oracle(K, V, n_docs, n_per, P):
    # K <= 32: warp-scan path (with and without the shared corpus copy); K = 40: sequential path
    doc, word = corpus(K, V, n_docs, n_per)
    n_tok = len(doc)
    ts = vb.B200TraceSet(vb.LDAModel(K, V, n_docs, n_tok), P, seed=81)
    o = OracleLDA(K, V, n_docs, n_tok, 1.0, 2.0, P, seed=81)
    o.seq(2, 0, 0, 0)
    ts.observe(doc, word)
    o.observe(doc, word)
    s = ts.rng_step
    ts.incorporate()
    o.seq(0, 0, n_tok, s)
    for scope, block, mode, K_tr, n_steps in [("z", "one", -1, 3 * n_tok, 3 * n_tok), ("z", "each", -3, 2, 2 * n_tok),
                                              ("z", 1, 1, 5, 5), ("default", 2 + 3, 3, 4, 4)]:
        s = ts.rng_step
        nodes, ch = ts.gibbs(scope, block, K_tr, return_changes=True)
        ref = o.gibbs(s, n_steps, mode)
        a, z = state_of(ts, n_tok)
        assert np.array_equal(z, o.z), (scope, block)
        assert np.array_equal(ch, ref)
        assert np.all(nodes == 7.0)
    assert ch.sum() >= 0
    # count tables stayed consistent with z: an MH sweep over every site still matches the oracle
    s = ts.rng_step
    ts.resimulation_mh("default", "each", 1)
    o.mh(s, 2 + n_tok, -3)
    a, z = state_of(ts, n_tok)
    assert np.array_equal(z, o.z)
    np.testing.assert_allclose(a, o.alpha, rtol=1e-12)


def test_lda_gibbs_vs_lite_fixture():
    c = CH["c4_gibbs"]
    recs, K, V = c["particles"], c["K"], c["V"]
    doc, word = np.array(c["doc"], dtype=np.int32), np.array(c["word"], dtype=np.int32)
    n_tok = len(doc)
    ts = vb.B200TraceSet(vb.LDAModel(K, V, c["n_docs"], 16, c["a"], c["b"]), c["P"], seed=c["seed"])
    zs = lambda: np.stack([ts.read_site(2 + i) for i in range(n_tok)], axis=1).astype(int)
    ts.rng_step = c["steps"]["prior"]
    ts.reset_to_prior()
    ts.observe(doc, word)
    ts.rng_step = c["steps"]["tokens"]
    ts.incorporate()
    assert np.array_equal(zs(), [r["z_tokens"] for r in recs])
    ts.rng_step = c["steps"]["gibbs_one"]
    ts.primitive_infer(["gibbs", "z", "one", c["n_gibbs"]])
    assert np.array_equal(zs(), [r["z_one"] for r in recs])
    ts.rng_step = c["steps"]["gibbs_each"]
    ts.primitive_infer(["gibbs", "z", "each", 2])
    assert np.array_equal(zs(), [r["z_each"] for r in recs])


def test_lda_gibbs_default_scope_is_not_enumerable():
    doc, word = corpus(3, 5, 2, 4)
    ts = vb.B200TraceSet(vb.LDAModel(3, 5, 2, len(doc)), 8, seed=1)
    ts.observe(doc, word)
    ts.incorporate()
    for block in ["one", "each", 0, 1]:
        with pytest.raises(vb.VentureException):
            ts.primitive_infer(["gibbs", "default", block, 1])
    with pytest.raises(vb.VentureException):      # continuous sites only
        t2 = vb.B200TraceSet(vb.NormalNormalModel(), 8, seed=1)
        t2.primitive_infer(["gibbs", "default", "one", 1])


def test_lda_gibbs_exact_posterior_tiny():
    # K = 2, V = 2, one document of 3 tokens, hyper-parameters pinned: the posterior over the 8 assignments
    # is enumerated exactly from the collapsed joint (dirichlet.py:192-201) and compared by chi-square under the
    # reference's policy (test/stats.py:101-131)
    from stats_policy import report_known_counts, statistical_test
    K, V, P = 2, 2, 40000
    doc, word = np.zeros(3, dtype=np.int32), np.array([0, 1, 0], dtype=np.int32)
    ad, aw = 0.7, 0.4
    lg = math.lgamma

    def logjoint(zz):
        ndk = np.bincount(zz, minlength=K)
        out = lg(K * ad) - lg(3 + K * ad) + sum(lg(ad + c) - lg(ad) for c in ndk)
        for k in range(K):
            cw = np.bincount(word[np.array(zz) == k], minlength=V)
            out += lg(V * aw) - lg(cw.sum() + V * aw) + sum(lg(aw + c) - lg(aw) for c in cw)
        return out
    confs = list(itertools.product(range(K), repeat=3))
    pr = np.exp([logjoint(np.array(c)) for c in confs])
    pr /= pr.sum()

    @statistical_test
    def run(seed):
        ts = vb.B200TraceSet(vb.LDAModel(K, V, 1, 3), P, seed=seed)
        ts.observe(doc, word)
        ts.incorporate(return_increments=False)
        ts.write_site(0, np.full(P, ad))
        ts.write_site(1, np.full(P, aw))
        ts.refresh()
        ts.primitive_infer(["gibbs", "z", "each", 25])
        z = np.stack([ts.read_site(2 + i) for i in range(3)], axis=1).astype(int)
        obs = np.array([(z == np.array(c)).all(1).sum() for c in confs])
        return report_known_counts(pr, obs)
    run()
