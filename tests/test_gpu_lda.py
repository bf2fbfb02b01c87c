"""GPU parity for config C4 (examples/lda.vnt, collapsed sym_dir_cat counts)
against the oracle: sequential observe/incorporate, likelihood_weight, z-site
and alpha-site (AAA) resimulation MH, resampling of the count tables."""
import numpy as np
import pytest

import venturecxx_b200 as vb
from helpers import OracleLDA

pytestmark = pytest.mark.gpu


def corpus(K, V, n_docs, n_per, seed=3):
    rng = np.random.default_rng(seed)
    theta = rng.dirichlet(np.full(K, 0.5), n_docs)
    phi = rng.dirichlet(np.full(V, 0.5), K)
    doc, word = [], []
    for d in range(n_docs):
        z = rng.choice(K, n_per, p=theta[d])
        for k in z:
            doc.append(d); word.append(rng.choice(V, p=phi[k]))
    return np.array(doc, dtype=np.int32), np.array(word, dtype=np.int32)


def state_of(ts, n_tok):
    alpha = np.stack([ts.read_site(0), ts.read_site(1)], axis=1)
    z = np.stack([ts.read_site(2 + i) for i in range(n_tok)], axis=1).astype(np.uint8) if n_tok else None
    return alpha, z


@pytest.mark.parametrize("K,V,n_docs,n_per,P", [(2, 2, 3, 4, 200), (5, 40, 4, 16, 150), (40, 7, 2, 9, 64)])
def test_lda_chain_matches_oracle(K, V, n_docs, n_per, P):
    doc, word = corpus(K, V, n_docs, n_per)
    n_tok = len(doc)
    ts = vb.B200TraceSet(vb.LDAModel(K, V, n_docs, n_tok), P, seed=51)
    o = OracleLDA(K, V, n_docs, n_tok, 1.0, 2.0, P, seed=51)
    o.seq(2, 0, 0, 0)            # the alphas' `assume` simulation (reset_to_prior at rng step 0)
    a, _ = state_of(ts, 0)
    np.testing.assert_allclose(a, o.alpha, rtol=1e-12)
    # observe in two batches, incorporating each (engine.observe incorporates immediately)
    h = n_tok // 2
    for lo, hi in [(0, h), (h, n_tok)]:
        ts.observe(doc[lo:hi], word[lo:hi])
        o.observe(doc[lo:hi], word[lo:hi])
        s = ts.rng_step
        inc = ts.incorporate()
        np.testing.assert_allclose(inc, o.seq(0, lo, hi, s), rtol=1e-12)
    a, z = state_of(ts, n_tok)
    assert np.array_equal(z, o.z)
    np.testing.assert_allclose(ts.log_weights, o.lw, rtol=1e-12)
    assert ts.numBlocksInScope() == 2 + n_tok
    # MH: random sites (z and alpha), then one full ordered sweep, then the two AAA sites alone
    for block, mode, K_tr, n_steps in [("one", -1, 3 * (2 + n_tok), 3 * (2 + n_tok)), ("each", -3, 1, 2 + n_tok),
                                       (0, 0, 6, 6), (1, 1, 6, 6)]:
        s = ts.rng_step
        _, acc = ts.resimulation_mh("default", block, K_tr, return_accepts=True)
        ref_acc = o.mh(s, n_steps, mode)
        a, z = state_of(ts, n_tok)
        assert np.array_equal(z, o.z), block
        np.testing.assert_allclose(a, o.alpha, rtol=1e-12)
        assert np.array_equal(acc, ref_acc), block
    assert acc.sum() > 0
    # likelihood_weight resimulates everything from the prior
    s = ts.rng_step
    ts.likelihood_weight()
    o.seq(1, 0, n_tok, s)
    a, z = state_of(ts, n_tok)
    assert np.array_equal(z, o.z)
    np.testing.assert_allclose(ts.log_weights, o.lw, rtol=1e-12)
    # resample moves alpha, z and every count table together: a later sweep still matches
    anc = ts.resample(P, mode="systematic", return_ancestors=True)
    for arr in ["alpha", "nk", "ndk", "nkw", "z"]:
        setattr(o, arr, np.ascontiguousarray(getattr(o, arr)[anc]))
    s = ts.rng_step
    ts.resimulation_mh("default", "one", 50)
    o.mh(s, 50, -1)
    a, z = state_of(ts, n_tok)
    assert np.array_equal(z, o.z)
    np.testing.assert_allclose(a, o.alpha, rtol=1e-12)


def test_write_site_refresh_recounts():
    K, V, n_docs, n_per, P = 3, 6, 2, 5, 32
    doc, word = corpus(K, V, n_docs, n_per)
    n_tok = len(doc)
    ts = vb.B200TraceSet(vb.LDAModel(K, V, n_docs, n_tok), P, seed=9)
    ts.observe(doc, word)
    ts.incorporate()
    o = OracleLDA(K, V, n_docs, n_tok, 1.0, 2.0, P, seed=9)
    o.observe(doc, word)
    o.seq(2, 0, 0, 0)
    o.seq(0, 0, n_tok, 1)
    # inject a different assignment for token 0 in every particle, recount, and compare a z-site weight
    znew = (o.z[:, 0].astype(int) + 1) % K
    ts.write_site(2, znew.astype(float))
    ts.refresh()
    for p in range(P):
        zo = o.z[p, 0]
        o.nkw[p, zo * V + word[0]] -= 1; o.nk[p, zo] -= 1; o.ndk[p, doc[0] * K + zo] -= 1
        zn = znew[p]; o.z[p, 0] = zn
        o.nkw[p, zn * V + word[0]] += 1; o.nk[p, zn] += 1; o.ndk[p, doc[0] * K + zn] += 1
    got = ts.log_likelihood_at("default", 2)
    aw = o.alpha[:, 1]
    ref = np.log((o.nkw[np.arange(P), znew * V + word[0]] - 1 + aw) / (o.nk[np.arange(P), znew] - 1 + aw * V))
    np.testing.assert_allclose(got, ref, rtol=1e-12)
