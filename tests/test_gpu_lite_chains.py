"""The CUDA path, through the C ABI, against chains produced by the reference's own Lite backend.

tests/golden/lite_chains.json holds the states and weights that backend/lite (Trace, constructScaffold,
detachAndExtract, regenAndAttach, mixMH, MHOperator, subsampledMixMH, makeConsistent, likelihood_weight)
ended with when fed the draws of Philox stream (seed, particle, step) -- see
tests/golden/make_lite_chains.py.  The kernels replay the same (seed, particle, step):
fp64 values to 1e-12 relative, discrete state exactly, transloop's node counts exactly."""
import numpy as np
import pytest

import venturecxx_b200 as vb
from helpers import golden, unj

pytestmark = pytest.mark.gpu
RTOL = 1e-12
CH = golden("lite_chains.json")


def col(recs, key):
    return np.array([unj(r[key]) for r in recs], dtype=np.float64)


def wcol(recs, key):
    return np.array([unj(r[key]) for r in recs], dtype=np.float64).T


def test_c1_normal_normal_vs_lite():
    c = CH["c1"]
    recs, y, K = c["particles"], c["y"], c["K"]
    ts = vb.B200TraceSet(vb.NormalNormalModel(c["mu0"], c["s0"], c["s"]), c["P"], seed=c["seed"])
    ts.observe(y[:2])
    ts.incorporate()
    ts.rng_step = c["steps"]["likelihood_weight"]
    ts.likelihood_weight()
    np.testing.assert_allclose(ts.read_site(0), col(recs, "x_lw"), rtol=RTOL)
    np.testing.assert_allclose(ts.log_weights, col(recs, "lw"), rtol=RTOL)
    ts.observe(y[2:])
    np.testing.assert_allclose(ts.incorporate(), col(recs, "incr"), rtol=RTOL)
    ts.rng_step = c["steps"]["mh"]
    nodes = ts.primitive_infer(["resimulation_mh", "default", "one", K])
    np.testing.assert_allclose(ts.read_site(0), col(recs, "x_mh"), rtol=RTOL)
    assert np.array_equal(nodes, [np.mean(r["nodes"]) for r in recs])
    np.testing.assert_allclose(ts.log_likelihood_at("default", "all"), col(recs, "ll_at"), rtol=RTOL)
    np.testing.assert_allclose(ts.log_joint_at("default", "all"), col(recs, "lj_at"), rtol=RTOL)


def test_c2_logistic_regression_vs_lite():
    c = CH["c2"]
    recs, D, N, K = c["particles"], c["D"], c["N"], c["K"]
    X, y = np.array(c["X"]), np.array(c["y"], dtype=np.uint8)
    h = N // 2
    ts = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), c["P"], seed=c["seed"])
    rw = lambda: np.stack([ts.read_site(j) for j in range(D)])
    ts.observe(X[:h], y[:h])
    ts.incorporate()
    ts.rng_step = c["steps"]["likelihood_weight"]
    ts.likelihood_weight()
    np.testing.assert_allclose(rw(), wcol(recs, "w_lw"), rtol=RTOL)
    np.testing.assert_allclose(ts.log_weights, col(recs, "lw"), rtol=RTOL)
    ts.observe(X[h:], y[h:])
    np.testing.assert_allclose(ts.incorporate(), col(recs, "incr"), rtol=RTOL)
    ts.rng_step = c["steps"]["mh_one"]
    nodes = ts.primitive_infer(["resimulation_mh", "default", "one", K])
    np.testing.assert_allclose(rw(), wcol(recs, "w_one"), rtol=RTOL)
    np.testing.assert_allclose(ts.log_likelihood_at("default", "all"), col(recs, "ll_one"), rtol=RTOL)
    assert np.all(nodes == c["untagged_nodes_one"][0])          # 3 + 2N (the model without the tag wrapper)
    ts.rng_step = c["steps"]["mh_each"]
    ts.primitive_infer(["resimulation_mh", "default", "each", 2])
    np.testing.assert_allclose(rw(), wcol(recs, "w_each"), rtol=RTOL)
    ts.rng_step = c["steps"]["mh_all"]
    nodes = ts.primitive_infer(["resimulation_mh", "default", "all", 4])
    np.testing.assert_allclose(rw(), wcol(recs, "w_all"), rtol=RTOL)
    assert np.all(nodes == c["untagged_nodes_all"][0])          # 2D + 1 + 2N
    np.testing.assert_allclose(ts.log_joint_at("default", "all"), col(recs, "lj_all"), rtol=1e-11)


def test_c3_hmm_vs_lite():
    c = CH["c3"]
    recs, T, K = c["particles"], c["T"], c["K"]
    ts = vb.B200TraceSet(vb.HMMModel(c["trans"], c["emit"], 1, 32), c["P"], seed=c["seed"])
    hist = lambda: np.stack([ts.read_site(t) for t in range(T)], axis=1).astype(int)
    for t in range(T):
        ts.observe(c["data"][t])
        ts.rng_step = c["steps"]["filter"] + t
        np.testing.assert_allclose(ts.incorporate(), [unj(r["incr"][t]) for r in recs], rtol=RTOL)
    assert np.array_equal(hist(), [r["hist_filter"] for r in recs])
    ts.rng_step = c["steps"]["mh_one"]
    nodes = ts.primitive_infer(["resimulation_mh", "default", "one", K])
    assert np.array_equal(hist(), [r["hist_one"] for r in recs])
    np.testing.assert_allclose(nodes, [np.mean(r["nodes_one"]) for r in recs], rtol=1e-15)
    ts.rng_step = c["steps"]["mh_each"]
    nodes = ts.primitive_infer(["resimulation_mh", "default", "each", 2])
    assert np.array_equal(hist(), [r["hist_each"] for r in recs])
    # `each`: scaffolder_loop sums the blocks' counts and floor-divides by #blocks (dispatch.py:122-128)
    per_sweep = [sum(r["nodes_each"][:T - 1]) // (T - 1) for r in recs]
    np.testing.assert_allclose(nodes, per_sweep, rtol=1e-15)
    np.testing.assert_allclose(ts.log_likelihood_at("default", "all"), col(recs, "ll_at"), rtol=1e-11)
    np.testing.assert_allclose(ts.log_joint_at("default", "all"), col(recs, "lj_at"), rtol=1e-11)


def test_c4_lda_vs_lite():
    c = CH["c4"]
    recs, K, V = c["particles"], c["K"], c["V"]
    doc, word = np.array(c["doc"], dtype=np.int32), np.array(c["word"], dtype=np.int32)
    n_tok = len(doc)
    ts = vb.B200TraceSet(vb.LDAModel(K, V, c["n_docs"], 16, c["a"], c["b"]), c["P"], seed=c["seed"])
    alpha = lambda: np.stack([ts.read_site(0), ts.read_site(1)], axis=1)
    zs = lambda: np.stack([ts.read_site(2 + i) for i in range(n_tok)], axis=1).astype(int)
    ts.rng_step = c["steps"]["prior"]
    ts.reset_to_prior()
    np.testing.assert_allclose(alpha(), [unj(r["alpha_prior"]) for r in recs], rtol=RTOL)
    ts.observe(doc, word)
    ts.rng_step = c["steps"]["tokens"]
    np.testing.assert_allclose(ts.incorporate(), [sum(unj(r["incr"])) for r in recs], rtol=RTOL)
    assert np.array_equal(zs(), [r["after_tokens"]["z"] for r in recs])
    ts.rng_step = c["steps"]["mh_one"]
    nodes = ts.primitive_infer(["resimulation_mh", "default", "one", c["n_mh"]])
    assert np.array_equal(zs(), [r["after_one"]["z"] for r in recs])
    np.testing.assert_allclose(alpha(), [unj(r["after_one"]["alpha"]) for r in recs], rtol=RTOL)
    # z site: 7 nodes; alpha_document_topic: 1 + 2 n_docs; alpha_topic_word: 1 + 2 (topics in use)
    np.testing.assert_allclose(nodes, [np.mean(r["nodes_one"]) for r in recs], rtol=1e-15)


def test_c5_subsampled_mh_vs_lite():
    c = CH["c5"]
    recs, D, N, K = c["particles"], c["D"], c["N"], c["K"]
    X, y = np.array(c["X"]), np.array(c["y"], dtype=np.uint8)
    ts = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), c["P"], seed=c["seed"])   # prior at rng step 0
    rw = lambda: np.stack([ts.read_site(j) for j in range(D)])
    np.testing.assert_allclose(rw(), wcol(recs, "w_prior"), rtol=RTOL)
    ts.observe(X, y)
    ts.incorporate()
    for k in range(K):
        ts.rng_step = c["steps"]["ssmh"] + k
        st = ts.primitive_infer(["subsampled_mh", "default", "one", c["Nbatch"], c["k0"], c["eps"], False, 0, False, 1])
        assert np.array_equal(st[:, 0].astype(bool), [r["accept"][k] for r in recs]), k
        assert np.array_equal(st[:, 1], [r["n"][k] for r in recs]), k
        np.testing.assert_allclose(st[:, 2], [unj(r["tstat"][k]) for r in recs], rtol=1e-8)
    np.testing.assert_allclose(rw(), wcol(recs, "w"), rtol=RTOL)


def test_resample_parents_vs_lite():
    """trace_set.py:202-219 on the reference's own sampleLogCategorical / log_domain_even_out."""
    for name, c in CH["resample"].items():
        lw = np.array(unj(c["lw"]), dtype=np.float64)
        P = len(lw)
        ts = vb.B200TraceSet(vb.NormalNormalModel(), P, seed=1)
        ts.set_particle_log_weights(lw)
        x0 = ts.read_site(0)
        anc = ts.resample(P, mode="multinomial", uniforms=np.array(c["uniforms"]), return_ancestors=True)
        assert np.array_equal(anc, c["parents"]), name
        assert np.array_equal(ts.read_site(0), x0[anc]), name          # state moved with the parent
        new = np.array(unj(c["new_lw"]), dtype=np.float64)
        got = ts.log_weights
        assert np.array_equal(np.isneginf(got), np.isneginf(new)), name
        fin = np.isfinite(new)
        np.testing.assert_allclose(got[fin], new[fin], rtol=RTOL)
