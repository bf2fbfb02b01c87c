"""Device particles <-> Lite traces.  A particle of a B200TraceSet, after inference on the
GPU, is dumped in the reference's serialisation (B200TraceSet.dump_particle), restored into a real Lite
trace by engine.trace.Trace.restore, and Lite's log_joint_at of that trace must equal the device's
log_joint_at of the particle at 1e-12: "identical traces" made literal.  The other way, dumps of Lite traces
are loaded into the device particles (B200TraceSet.load)."""
import numpy as np
import pytest

import venturecxx_b200 as vb
from helpers import golden
from test_interchange_lite import (REF, collapsed_lda_log_joint, lite, lite_dump, log_joint, restore,
                                   serialized_of)
import os

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "venture", "lite")),
                                 reason="oracle/_ref (py3 rendering of Lite) not built")]
H = golden("hmm.json")


def sites_of(ts):
    return np.stack([ts.read_site(j) for j in range(ts.num_sites())])


def test_blr_particles_restore_into_lite():
    N, D, P = 40, 5, 24
    rng = np.random.default_rng(3)
    X = rng.normal(size=(N, D))
    y = (rng.random(N) < 0.5).astype(np.uint8)
    ts = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), P, seed=5)
    ts.observe(X[:25], y[:25])
    ts.incorporate()
    ts.observe(X[25:], y[25:])
    ts.incorporate()
    ts.resimulation_mh("default", "one", 50)
    lj = ts.log_joint_at("default", "all")
    w = sites_of(ts)
    for i in (0, 7, P - 1):
        et = restore(ts.dump_particle(i))
        assert [et.report_value(j + 1)["value"] for j in range(D)] == list(w[:, i])
        np.testing.assert_allclose(log_joint(et), lj[i], rtol=1e-12)
    assert len(ts.dump()) == P


def test_hmm_particles_restore_into_lite_after_filtering():
    P, T = 16, 8
    ts = vb.B200TraceSet(vb.HMMModel(H["transition_matrix"], H["emission_matrix"], 1, 32), P, seed=9)
    for d in H["data"][:T]:
        ts.observe(d)
        ts.incorporate()
        ts.resample(P, mode="systematic")
    ts.resimulation_mh("default", "one", 20)
    lj = ts.log_joint_at("default", "all")
    hist = sites_of(ts)
    for i in (0, 5, P - 1):
        ser = ts.dump_particle(i)
        et = restore(ser)
        np.testing.assert_allclose(log_joint(et), lj[i], rtol=1e-12)
        assert lite_dump(et)[0] == ser[0]
        assert list(hist[:, i]) == [1.0] + [v["value"] for v in list(reversed(ser[0]))[1::2]]


def test_lda_particles_restore_into_lite():
    K, V, n_docs, P = 3, 6, 3, 8
    rng = np.random.default_rng(4)
    doc = np.repeat(np.arange(n_docs), 4).astype(np.int32)
    word = rng.integers(0, V, len(doc)).astype(np.int32)
    model = vb.LDAModel(K, V, n_docs, len(doc))
    ts = vb.B200TraceSet(model, P, seed=13)
    ts.observe(doc, word)
    ts.incorporate()
    ts.resimulation_mh("default", "one", 60)
    s = sites_of(ts)
    c = {"K": K, "V": V, "a": 1.0, "b": 2.0, "n_docs": n_docs, "doc": doc.tolist(), "word": word.tolist()}
    for i in (0, 3, P - 1):
        et = restore(ts.dump_particle(i))
        np.testing.assert_allclose(log_joint(et), collapsed_lda_log_joint(c, s[:, i]), rtol=1e-12)


def test_lite_traces_load_into_device_particles():
    # P Lite traces with their own values -> device; the device's log joint of each equals Lite's
    N, D, P = 30, 4, 6
    rng = np.random.default_rng(8)
    X = rng.normal(size=(N, D))
    y = (rng.random(N) < 0.4).astype(np.uint8)
    model = vb.LogisticRegressionModel(D, N)
    ts = vb.B200TraceSet(model, P, seed=21)
    ts.observe(X, y)
    ts.incorporate()
    W = rng.normal(size=(D, P))
    dumps, lj_lite = [], []
    for i in range(P):
        et = restore(serialized_of(model, ts.observations(), W[:, i]))     # a Lite trace holding W[:, i]
        dumps.append(lite_dump(et))                                        # ... dumped by Lite itself
        lj_lite.append(log_joint(et))
    ts.load(dumps)
    np.testing.assert_array_equal(sites_of(ts), W)
    np.testing.assert_allclose(ts.log_joint_at("default", "all"), lj_lite, rtol=1e-12)
    with pytest.raises(vb.VentureException):
        ts.load(dumps[:-1])
    bad = list(dumps)
    bad[2] = (dumps[2][0], dict(list(dumps[2][1].items())[:-1]), set())
    with pytest.raises(vb.VentureException):
        ts.load(bad)


def test_normal_normal_dump_and_forget():
    ts = vb.B200TraceSet(vb.NormalNormalModel(10.0, 1.0, 1.0), 4, seed=2)
    ts.observe(14.0)
    ts.incorporate()
    ser = ts.dump_particle(1)
    et = restore(ser)
    assert et.report_value(1)["value"] == ts.read_site(0)[1]
    ts.forget_observations()
    assert ts.observations() == []
    assert len(ts.dump_particle(1)[1]) == 1          # only the assume is left
