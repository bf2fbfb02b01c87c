"""The oracle (oracle/oracle.c) against chains produced by the reference's own Lite backend.

tests/golden/lite_chains.json was written by tests/golden/make_lite_chains.py, which runs
backend/lite (Trace, constructScaffold, detachAndExtract, regenAndAttach, mixMH, MHOperator, ...)
with stand-in RNG objects that hand Lite the draws of Philox stream (seed, particle, step).  Here the
oracle replays the same (seed, particle, step) and must end in the same state with the same weights:
1e-12 relative for reals, exact for discrete state and for Lite's returned node counts."""
import numpy as np
import pytest

import oracle
from helpers import golden, unj

RTOL = 1e-12
CH = golden("lite_chains.json")


def col(recs, key):
    return np.array([unj(r[key]) for r in recs], dtype=np.float64)


# ------------------------------------------------------------------ C1
def test_c1_normal_normal_chain():
    c = CH["c1"]
    P, K, seed = c["P"], c["K"], c["seed"]
    mu0, s0, s = c["mu0"], c["s0"], c["s"]
    y = np.array(c["y"])
    recs = c["particles"]
    L = oracle.lib()
    x, lw = np.empty(P), np.empty(P)
    y2 = np.ascontiguousarray(y[:2])
    L.orc_nn_likelihood_weight(oracle._ptr(x), oracle._ptr(lw), oracle._ptr(y2), oracle.I64(2), oracle.I64(P),
                               oracle.I64(0), oracle.F64(mu0), oracle.F64(s0), oracle.F64(s), oracle.U64(seed),
                               oracle.U32(c["steps"]["likelihood_weight"]))
    np.testing.assert_allclose(x, col(recs, "x_lw"), rtol=RTOL)
    np.testing.assert_allclose(lw, col(recs, "lw"), rtol=RTOL)
    incr = np.empty(P)
    L.orc_nn_make_consistent(oracle._ptr(x), oracle._ptr(lw), oracle._ptr(incr), oracle._ptr(y), oracle.I64(2),
                             oracle.I64(3), oracle.I64(P), oracle.F64(s))
    np.testing.assert_allclose(incr, col(recs, "incr"), rtol=RTOL)
    acc = np.empty(P, dtype=np.int64)
    L.orc_nn_mh(oracle._ptr(x), oracle._ptr(y), oracle.I64(3), oracle.I64(P), oracle.I64(0), oracle.F64(mu0),
                oracle.F64(s0), oracle.F64(s), oracle.U64(seed), oracle.U32(c["steps"]["mh"]), oracle.I32(K),
                oracle._ptr(acc))
    np.testing.assert_allclose(x, col(recs, "x_mh"), rtol=RTOL)
    assert acc.sum() > 0 and (acc < K).any()          # the chain both accepted and rejected
    for joint, key in [(0, "ll_at"), (1, "lj_at")]:
        out = np.empty(P)
        L.orc_nn_log_likelihood_at(oracle._ptr(x), oracle._ptr(y), oracle.I64(3), oracle.I64(P), oracle.F64(mu0),
                                   oracle.F64(s0), oracle.F64(s), oracle.I32(joint), oracle._ptr(out))
        np.testing.assert_allclose(out, col(recs, key), rtol=RTOL)
    # transloop's return value (dispatch.py:109-120): |regenCounts| = x + one absorbing node per observation
    for r in recs:
        assert r["nodes"] == [1.0 + len(y)] * K


# ------------------------------------------------------------------ C2
def _blr_mh(w, X, y, seed, step0, n, mode):
    D, P = w.shape
    acc, ll = np.empty(P, dtype=np.int64), np.empty(P)
    oracle.lib().orc_blr_mh(oracle._ptr(w), oracle._ptr(X), oracle._ptr(y), oracle.I32(D), oracle.I64(X.shape[0]),
                            oracle.I64(P), oracle.I64(0), oracle.F64(0.0), oracle.F64(1.0), oracle.U64(seed),
                            oracle.U32(step0), oracle.I32(n), oracle.I32(mode), oracle._ptr(acc), oracle._ptr(ll))
    return acc, ll


def _blr_ll(w, X, y, i0, i1):
    D, P = w.shape
    out = np.empty(P)
    oracle.lib().orc_blr_log_likelihood(oracle._ptr(w), oracle._ptr(X), oracle._ptr(y), oracle.I32(D), oracle.I64(i0),
                                        oracle.I64(i1), oracle.I64(P), oracle._ptr(out))
    return out


def wcol(recs, key):
    return np.ascontiguousarray(np.array([unj(r[key]) for r in recs], dtype=np.float64).T)   # [D][P]


def test_c2_logistic_regression_chain():
    c = CH["c2"]
    P, D, N, K, seed = c["P"], c["D"], c["N"], c["K"], c["seed"]
    X = np.ascontiguousarray(c["X"], dtype=np.float64)
    y = np.ascontiguousarray(c["y"], dtype=np.uint8)
    recs = c["particles"]
    h = N // 2
    w = np.empty((D, P))
    oracle.lib().orc_blr_prior(oracle._ptr(w), oracle.I32(D), oracle.I64(P), oracle.I64(0), oracle.F64(0.0),
                               oracle.F64(1.0), oracle.U64(seed), oracle.U32(c["steps"]["likelihood_weight"]))
    np.testing.assert_allclose(w, wcol(recs, "w_lw"), rtol=RTOL)
    np.testing.assert_allclose(_blr_ll(w, X, y, 0, h), col(recs, "lw"), rtol=RTOL)
    np.testing.assert_allclose(_blr_ll(w, X, y, h, N), col(recs, "incr"), rtol=RTOL)
    acc, ll = _blr_mh(w, X, y, seed, c["steps"]["mh_one"], K, -1)
    np.testing.assert_allclose(w, wcol(recs, "w_one"), rtol=RTOL)
    np.testing.assert_allclose(ll, col(recs, "ll_one"), rtol=RTOL)
    assert acc.sum() > 0 and (acc < K).any()
    acc, _ = _blr_mh(w, X, y, seed, c["steps"]["mh_each"], 2 * D, -3)
    np.testing.assert_allclose(w, wcol(recs, "w_each"), rtol=RTOL)
    acc, ll = _blr_mh(w, X, y, seed, c["steps"]["mh_all"], 4, -2)
    np.testing.assert_allclose(w, wcol(recs, "w_all"), rtol=RTOL)
    prior = (-0.5 * np.log(2 * np.pi) - 0.5 * w * w).sum(0)
    np.testing.assert_allclose(ll + prior, col(recs, "lj_all"), rtol=1e-11)
    print(recs[0]["nodes_one"][0], recs[0]["nodes_each"][0], recs[0]["nodes_all"][0])


# ------------------------------------------------------------------ C3
def test_c3_hmm_filter_and_mh_chain():
    from helpers import OracleHMM
    c = CH["c3"]
    P, K, T, seed = c["P"], c["K"], c["T"], c["seed"]
    recs = c["particles"]
    o = OracleHMM(c["trans"], c["emit"], 1, P, 32, seed)
    for t in range(T):
        incr = o.step(c["data"][t], c["steps"]["filter"] + t)
        np.testing.assert_allclose(incr, [unj(r["incr"][t]) for r in recs], rtol=RTOL)
    assert np.array_equal(o.hist[:, :T], np.array([r["hist_filter"] for r in recs]))
    acc = o.mh(c["steps"]["mh_one"], K, -1)
    assert np.array_equal(o.hist[:, :T], np.array([r["hist_one"] for r in recs]))
    assert acc.sum() > 0 and (acc < K).any()
    o.mh(c["steps"]["mh_each"], 2 * (T - 1), -3)
    assert np.array_equal(o.hist[:, :T], np.array([r["hist_each"] for r in recs]))
    # log_likelihood_at(default, all) (dispatch.py:331-346): the absorbing border of the scaffold whose
    # principal nodes are the random choices state(2..T) -- the emission of the deterministic state(1)
    # is NOT a child of any principal node and does not count; log_joint_at adds the transitions.
    emit, trans = np.array(c["emit"]), np.array(c["trans"])
    ll = np.array([sum(np.log(emit[h[t] - 1, c["data"][t] - 1]) for t in range(1, T)) for h in o.hist])
    lp = np.array([sum(np.log(trans[h[t - 1] - 1, h[t] - 1]) for t in range(1, T)) for h in o.hist])
    np.testing.assert_allclose(ll, col(recs, "ll_at"), rtol=1e-11)
    np.testing.assert_allclose(ll + lp, col(recs, "lj_at"), rtol=1e-11)
    # transloop's return value: 13 affected nodes for an interior state(t), 9 for the last one
    for r in recs:
        assert set(r["nodes_each"]) == {13.0, 9.0} and r["nodes_each"][T - 2] == 9.0
    # enumerative Gibbs (infer/egibbs.py) continues the same chain
    ch = o.gibbs(c["steps"]["gibbs_one"], K, -1)
    assert np.array_equal(o.hist[:, :T], np.array([r["hist_gibbs_one"] for r in recs]))
    assert ch.sum() > 0
    o.gibbs(c["steps"]["gibbs_each"], 2 * (T - 1), -3)
    assert np.array_equal(o.hist[:, :T], np.array([r["hist_gibbs_each"] for r in recs]))
    for r in recs:
        assert set(r["nodes_gibbs_each"]) == {13.0, 9.0}


# ------------------------------------------------------------------ C4
def test_c4_lda_chain():
    from helpers import OracleLDA
    c = CH["c4"]
    P, K, V, seed = c["P"], c["K"], c["V"], c["seed"]
    doc, word = c["doc"], c["word"]
    n_tok = len(doc)
    recs = c["particles"]
    o = OracleLDA(K, V, c["n_docs"], 16, c["a"], c["b"], P, seed)
    o.seq(2, 0, 0, c["steps"]["prior"])                      # reset_to_prior: the two gamma hyper-parameters
    np.testing.assert_allclose(o.alpha, [unj(r["alpha_prior"]) for r in recs], rtol=RTOL)
    o.observe(doc, word)
    incr = o.seq(0, 0, n_tok, c["steps"]["tokens"])          # observe + incorporate token by token
    np.testing.assert_allclose(incr, [sum(unj(r["incr"])) for r in recs], rtol=RTOL)
    assert np.array_equal(o.z[:, :n_tok], np.array([r["after_tokens"]["z"] for r in recs]))
    acc = o.mh(c["steps"]["mh_one"], c["n_mh"], -1)
    assert np.array_equal(o.z[:, :n_tok], np.array([r["after_one"]["z"] for r in recs]))
    np.testing.assert_allclose(o.alpha, [unj(r["after_one"]["alpha"]) for r in recs], rtol=RTOL)
    assert acc.sum() > 0 and (acc < c["n_mh"]).any()
    # both hyper-parameter sites moved somewhere (the AAA weight was exercised with accepts)
    a0 = np.array([unj(r["alpha_prior"]) for r in recs])
    assert (o.alpha[:, 0] != a0[:, 0]).any() and (o.alpha[:, 1] != a0[:, 1]).any()


# ------------------------------------------------------------------ C5
def test_c5_subsampled_mh_chain():
    c = CH["c5"]
    P, D, N, K, seed = c["P"], c["D"], c["N"], c["K"], c["seed"]
    X = np.ascontiguousarray(c["X"], dtype=np.float64)
    y = np.ascontiguousarray(c["y"], dtype=np.uint8)
    recs = c["particles"]
    w = np.empty((D, P))
    oracle.lib().orc_blr_prior(oracle._ptr(w), oracle.I32(D), oracle.I64(P), oracle.I64(0), oracle.F64(0.0),
                               oracle.F64(1.0), oracle.U64(seed), oracle.U32(c["steps"]["prior"]))
    np.testing.assert_allclose(w, wcol(recs, "w_prior"), rtol=RTOL)
    stats = np.empty((P, 3))
    for k in range(K):
        oracle.lib().orc_blr_subsampled_mh(oracle._ptr(w), oracle._ptr(X), oracle._ptr(y), oracle.I32(D), oracle.I64(N),
                                           oracle.I64(P), oracle.I64(0), oracle.F64(0.0), oracle.F64(1.0),
                                           oracle.U64(seed), oracle.U32(c["steps"]["ssmh"] + k), oracle.I32(1),
                                           oracle.I32(-1), oracle.F64(c["Nbatch"]), oracle.I32(c["k0"]),
                                           oracle.F64(c["eps"]), oracle.I32(0), oracle.F64(0.0), oracle._ptr(stats))
        for p, r in enumerate(recs):
            assert bool(stats[p, 0]) == r["accept"][k], (k, p)
            assert stats[p, 1] == r["n"][k], (k, p)            # sections visited before the test stopped
            np.testing.assert_allclose(stats[p, 2], unj(r["tstat"][k]), rtol=1e-9)
    np.testing.assert_allclose(w, wcol(recs, "w"), rtol=RTOL)
    ns = np.array([r["n"] for r in recs])
    assert (ns < N).any() and (ns == N).any()      # early stops and full passes both occurred


def test_c4_lda_enumerative_gibbs_chain():
    from helpers import OracleLDA
    c = CH["c4_gibbs"]
    P, K, V, seed = c["P"], c["K"], c["V"], c["seed"]
    doc, word = c["doc"], c["word"]
    n_tok = len(doc)
    recs = c["particles"]
    o = OracleLDA(K, V, c["n_docs"], 16, c["a"], c["b"], P, seed)
    o.seq(2, 0, 0, c["steps"]["prior"])
    o.observe(doc, word)
    o.seq(0, 0, n_tok, c["steps"]["tokens"])
    assert np.array_equal(o.z[:, :n_tok], np.array([r["z_tokens"] for r in recs]))
    np.testing.assert_allclose(o.alpha, [unj(r["alpha"]) for r in recs], rtol=RTOL)
    ch = o.gibbs(c["steps"]["gibbs_one"], c["n_gibbs"], -1)
    assert np.array_equal(o.z[:, :n_tok], np.array([r["z_one"] for r in recs]))
    assert ch.sum() > 0
    o.gibbs(c["steps"]["gibbs_each"], 2 * n_tok, -3)
    assert np.array_equal(o.z[:, :n_tok], np.array([r["z_each"] for r in recs]))
    # the tagged z site's scaffold has 8 nodes (7 of lda.vnt's + the tag application)
    assert all(set(r["nodes_one"]) == {8.0} for r in recs)


# ------------------------------------------------------------------ resample
def test_resample_parents_and_weights_vs_lite():
    """trace_set.py:202-219: parents drawn by the reference's sampleLogCategorical at prescribed uniforms
    equal the oracle's multinomial ancestors (64-bit fixed-point CDF) at the same uniforms; the new
    weights equal log_domain_even_out."""
    for name, c in CH["resample"].items():
        lw = np.array(unj(c["lw"]), dtype=np.float64)
        anc = oracle.ancestors(lw, len(lw), 0, uniforms=np.array(c["uniforms"]))
        assert np.array_equal(anc, c["parents"]), name
        new = oracle.logsumexp(lw) - np.log(len(lw))
        for v in unj(c["new_lw"]):
            assert (np.isneginf(new) and np.isneginf(v)) or abs(new - v) <= 1e-12 * abs(v), name


# ------------------------------------------------------------------ provenance
def test_c1_fixture_regenerates_from_live_lite():
    """Where the py3 rendering of the reference's Lite backend exists (oracle/_ref, built from /root/reference by
    oracle/lite_py3/build_lite_py3.py -- this container, not the GPU box), run Lite again and demand the
    committed C1 chain bit for bit.  (All chains regenerate identically; the others take minutes.)"""
    import os
    import sys
    ref = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "venture")
    if not os.path.isdir(ref):
        pytest.skip("oracle/_ref not built here")
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
    import make_lite_chains as M
    import json
    fresh = json.loads(json.dumps(M.chain_c1()))
    assert fresh == CH["c1"]
