"""N4 (pgibbs): `(pgibbs seq ordered Pi K)` on the HMM of examples/hmm.vnt against the reference.

tests/golden/pgibbs_chains.json holds what Lite's PGibbsOperator (infer/pgibbs.py:49-126, under mixMH) did to real
Lite traces whose RNGs were the oracle's samplers on the Philox streams (tests/golden/make_pgibbs_chains.py): the
state sequence after every transition and transloop's node counts.
  CPU: oracle.c (orc_hmm_pgibbs) replays the same streams -- every state of every transition equal.
  GPU: vk_pgibbs does; and ordered particle-Gibbs sweeps reach the exact smoothed marginals of hmm_plugin.py.
"""
import numpy as np
import pytest

from helpers import OracleHMM, golden

FX = golden("pgibbs_chains.json")


@pytest.mark.parametrize("name", ["hmm", "hmm_one_particle"])
def test_oracle_pgibbs_equals_lite(name):
    c = FX[name]
    recs, T, K, Pi = c["particles"], c["T"], c["K"], c["Pi"]
    o = OracleHMM(c["trans"], c["emit"], 1, c["P"], 32, c["seed"])
    o.T, o.symbols = T, list(c["data"])
    o.hist[:, :T] = np.array([r["hist_start"] for r in recs], dtype=np.uint8)
    changed = 0
    for k in range(K):
        before = o.hist[:, :T].copy()
        acc = o.pgibbs(c["steps"]["pgibbs"] + k, 1, Pi)
        assert np.array_equal(o.hist[:, :T], np.array([r["hists"][k] for r in recs])), k
        changed += int((before != o.hist[:, :T]).any(axis=1).sum())
        assert np.all((acc == 0) | (acc == 1))
    assert changed > 0
    # one scaffold over all state sites: 13 nodes per interior state, 9 for the newest (scaffold.py:67-68)
    assert all(n == 13 * (T - 2) + 9 for r in recs for n in r["nodes"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["hmm", "hmm_one_particle"])
def test_device_pgibbs_equals_lite(name):
    import venturecxx_b200 as vb
    c = FX[name]
    recs, T, K, Pi = c["particles"], c["T"], c["K"], c["Pi"]
    ts = vb.B200TraceSet(vb.HMMModel(c["trans"], c["emit"], 1, 32), c["P"], seed=c["seed"])
    for d in c["data"]:
        ts.observe(d)
        ts.incorporate(return_increments=False)
    for t in range(1, T):
        ts.write_site(t, np.array([r["hist_start"][t] for r in recs], dtype=np.float64))
    for k in range(K):
        ts.rng_step = c["steps"]["pgibbs"] + k
        nodes = ts.primitive_infer(["pgibbs", "seq", "ordered", Pi, 1])
        hist = np.stack([ts.read_site(t) for t in range(T)], axis=1).astype(int)
        assert np.array_equal(hist, np.array([r["hists"][k] for r in recs])), k
        assert np.all(nodes == 13 * (T - 2) + 9)
    # K transitions in one call continue the same chains
    ts2 = vb.B200TraceSet(vb.HMMModel(c["trans"], c["emit"], 1, 32), c["P"], seed=c["seed"])
    for d in c["data"]:
        ts2.observe(d)
        ts2.incorporate(return_increments=False)
    for t in range(1, T):
        ts2.write_site(t, np.array([r["hist_start"][t] for r in recs], dtype=np.float64))
    ts2.rng_step = c["steps"]["pgibbs"]
    _, acc = ts2.pgibbs("seq", "ordered", Pi, K, return_accepts=True)
    assert np.array_equal(np.stack([ts2.read_site(t) for t in range(T)], axis=1).astype(int), np.array([r["hists"][K - 1] for r in recs]))
    assert acc.max() <= K
    with pytest.raises(vb.VentureException):
        ts2.primitive_infer(["pgibbs", "default", "one", Pi, 1])            # pgibbs wants an ordered scope
    with pytest.raises(vb.VentureException):
        vb.B200TraceSet(vb.NormalNormalModel(), 4).primitive_infer(["pgibbs", "default", "ordered", 4, 1])


@pytest.mark.gpu
def test_pgibbs_reaches_smoothed_marginals():
    """hmm.vnt's 20 observations, then particle-Gibbs transitions: the outer particles' state marginals approach the
    exact smoothed marginals tabulated in examples/hmm_plugin.py:130-173 (TVD, :225-232) -- far faster than single-site
    Gibbs, which is what the operator is for"""
    import venturecxx_b200 as vb
    H = golden("hmm.json")
    P, T = 1 << 14, 20
    ts = vb.B200TraceSet(vb.HMMModel(H["transition_matrix"], H["emission_matrix"], 1, 20), P, seed=7)
    for d in H["data"]:
        ts.observe(d)
        ts.incorporate(return_increments=False)
    _, acc = ts.pgibbs("seq", "ordered", 16, 150, return_accepts=True)
    assert acc.mean() > 3
    hist = np.stack([ts.read_site(t) for t in range(T)], axis=1).astype(int)
    sm = np.array(H["smoothed_marginals"])
    worst = 0.0
    for t in range(T):
        m = np.bincount(hist[:, t], minlength=6)[1:6] / float(P)
        worst = max(worst, 0.5 * np.abs(m - sm[t]).sum())
    assert worst < 0.03, worst
