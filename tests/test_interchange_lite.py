"""State interchange with Lite, checked with the reference's own code: the dump a compiled
model writes for a particle (venturecxx_b200/interchange.py) is restored by engine.trace.Trace.restore
(python/lib/engine/trace.py:168-171) into a real Lite trace, and

  * Lite's own dump of that trace (engine/trace.py:164-166, 179-202) is equal to ours, value for value and
    directive for directive -- order of the value stack and boxing included;
  * the restored trace reports the particle's values, and at the states recorded in
    tests/golden/lite_chains.json its log_joint_at equals what Lite computed when it produced them.

CPU only: the product side here is host code (program text + value order); tests/test_gpu_interchange.py
does the same from device particles.  Needs oracle/_ref (the py3 rendering of the reference, built by
__graft_entry__.build() where /root/reference exists)."""
import os
import sys
import warnings

import numpy as np
import pytest

import venturecxx_b200.models as M
from venturecxx_b200 import interchange as ix
from helpers import golden, unj

REF = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref")
pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "venture", "lite")),
                                reason="oracle/_ref (py3 rendering of Lite) not built")
CH = golden("lite_chains.json")


def lite():
    warnings.simplefilter("ignore")
    if REF not in sys.path:
        sys.path.insert(0, REF)
    from venture.engine.trace import Trace as ETrace
    from venture.lite.trace import Trace as LTrace
    import venture.lite.value as vv
    return ETrace, LTrace, vv


def restore(serialized):
    ETrace, LTrace, _ = lite()
    return ETrace.restore(lambda: LTrace(1), serialized, {})


def lite_dump(et):
    values, directives, names = et.dump()
    return list(values), directives, names          # py3: dumpSerializationDB returns a map object


def serialized_of(model, obs, sites):
    return (list(reversed(model.random_values(obs, np.asarray(sites, dtype=np.float64)))),
            ix.as_directive_dict(model.program(obs)), set())


def log_joint(et):
    _, _, vv = lite()
    return et.trace.log_joint_at(vv.VentureSymbol("default"), vv.VentureSymbol("all"))


def check_roundtrip(model, obs, sites):
    ser = serialized_of(model, obs, sites)
    et = restore(ser)
    back = lite_dump(et)
    assert back[0] == ser[0]
    assert ix._strip(back[1]) == ix._strip(ser[1])
    np.testing.assert_array_equal(model.sites_from_random_values(obs, list(reversed(back[0]))), sites)
    return et


def test_normal_normal_roundtrip_and_fixture_log_joint():
    c = CH["c1"]
    m = M.NormalNormalModel(c["mu0"], c["s0"], c["s"])
    obs = [(c["y"][:2],), (c["y"][2:],)]
    for r in c["particles"]:
        et = check_roundtrip(m, obs, [unj(r["x_mh"])])
        assert et.report_value(1)["value"] == unj(r["x_mh"])
        np.testing.assert_allclose(log_joint(et), unj(r["lj_at"]), rtol=1e-13)


def test_logistic_regression_roundtrip_and_fixture_log_joint():
    c = CH["c2"]
    D, N = c["D"], c["N"]
    X, y = np.array(c["X"]), np.array(c["y"])
    m = M.LogisticRegressionModel(D, N)
    obs = [(X[:N // 2], y[:N // 2]), (X[N // 2:], y[N // 2:])]
    for r in c["particles"]:
        w = np.array(unj(r["w_all"]), dtype=np.float64)
        et = check_roundtrip(m, obs, w)
        assert [et.report_value(j + 1)["value"] for j in range(D)] == list(w)
        np.testing.assert_allclose(log_joint(et), unj(r["lj_all"]), rtol=1e-12)


def test_hmm_roundtrip_and_fixture_log_joint():
    c = CH["c3"]
    m = M.HMMModel(c["trans"], c["emit"], 1, 32)
    obs = [(d,) for d in c["data"][:c["T"]]]
    for r in c["particles"]:
        et = check_roundtrip(m, obs, np.array(r["hist_each"], dtype=np.float64))
        np.testing.assert_allclose(log_joint(et), unj(r["lj_at"]), rtol=1e-12)


def collapsed_lda_log_joint(c, sites):
    """gamma hyper-priors + the two Dirichlet-multinomial marginals (dirichlet.py:192-201 logDensityOfData)."""
    import math
    lg = math.lgamma
    K, V, a, b = c["K"], c["V"], c["a"], c["b"]
    ad, aw = sites[0], sites[1]
    z, doc, word = np.array(sites[2:], dtype=int), np.array(c["doc"]), np.array(c["word"])

    def gam(x):
        return a * math.log(b) - lg(a) + (a - 1) * math.log(x) - b * x
    tot = gam(ad) + gam(aw)
    for d in range(c["n_docs"]):
        n = np.bincount(z[doc == d], minlength=K)
        if n.sum():
            tot += lg(K * ad) - lg(n.sum() + K * ad) + sum(lg(ad + x) - lg(ad) for x in n)
    for k in range(K):
        n = np.bincount(word[z == k], minlength=V)
        if n.sum():
            tot += lg(V * aw) - lg(n.sum() + V * aw) + sum(lg(aw + x) - lg(aw) for x in n)
    return tot


def test_lda_roundtrip_and_collapsed_joint():
    c = CH["c4"]
    m = M.LDAModel(c["K"], c["V"], c["n_docs"], len(c["doc"]), c["a"], c["b"])
    obs = [(np.array(c["doc"][:5]), np.array(c["word"][:5])), (np.array(c["doc"][5:]), np.array(c["word"][5:]))]
    for r in c["particles"]:
        st = r["after_one"]
        sites = np.array(list(unj(st["alpha"])) + list(st["z"]), dtype=np.float64)
        et = check_roundtrip(m, obs, sites)
        assert et.report_value(3)["value"] == sites[0] and et.report_value(4)["value"] == sites[1]
        # the replay rebuilt the collapsed makers' counts: Lite's joint of the restored trace is the
        # Dirichlet-multinomial joint of exactly these assignments
        np.testing.assert_allclose(log_joint(et), collapsed_lda_log_joint(c, sites), rtol=1e-12)


def test_load_is_the_inverse_of_dump_on_the_host_side():
    m = M.HMMModel(CH["c3"]["trans"], CH["c3"]["emit"], 1, 32)
    obs = [(d,) for d in CH["c3"]["data"][:6]]
    sites = np.array([1, 2, 3, 3, 4, 5], dtype=np.float64)
    ser = serialized_of(m, obs, sites)
    np.testing.assert_array_equal(m.sites_from_random_values(obs, list(reversed(ser[0]))), sites)
    # a dump of a different program is refused by content, not by identity
    other = serialized_of(m, obs[:5], sites[:5])
    assert ix._strip(other[1]) != ix._strip(ser[1])
