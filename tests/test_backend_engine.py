"""B1/B2/B4: the reference's own Engine (imported from oracle/_ref, the py3 rendering of the reference; test
infrastructure) driving the B200 backend through its stock seams -- `Backend.trace_constructor`, the Master
protocol, `TraceSet` -- with the reference's inference programs UNCHANGED.

  * CPU (`-m "not gpu"`): the seam logic (directive decoding, Master protocol, TraceSet overrides, inference
    SP routing) over a numpy stand-in for the device batch.
  * GPU: the same programs on the real batch; weights and chains equal the direct C-ABI calls bit for bit
    (same seed, same sequence of entry points).
"""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import ref_engine  # noqa: E402

pytestmark = pytest.mark.skipif(not ref_engine.available(), reason="oracle/_ref (py3 rendering of the reference) not built")


def _engine(backend, seed=11):
    ref_engine._import()
    ref_engine._install_prelude_reader()
    return backend.make_engine(seed=seed)


# ---- a numpy stand-in for the device batch (normal-normal model), CPU tests only ---------------------------------
class FakeBatch(object):
    """the B200TraceSet methods backend.py uses, for NormalNormalModel, in numpy"""

    def __init__(self, model, n, seed):
        self.model, self.n_particles, self.seed = model, n, seed
        self.rng = np.random.default_rng(seed)
        self.x = self.rng.normal(model.prior_mu, model.prior_sigma, size=n)
        self.lw = np.zeros(n)
        self.obs, self.n_prop = [], 0
        self.calls = []

    def _ll(self, ys):
        s = self.model.obs_sigma
        return sum(-np.log(s) - 0.5 * np.log(2 * np.pi) - 0.5 * ((y - self.x) / s) ** 2 for y in ys) + np.zeros(self.n_particles)

    def observe(self, y):
        self.obs.append((y,))

    def observations(self):
        return list(self.obs)

    def incorporate(self, return_increments=True):
        inc = self._ll([y for (y,) in self.obs[self.n_prop:]])
        self.n_prop = len(self.obs)
        self.lw = self.lw + inc
        return inc

    def likelihood_weight(self):
        self.calls.append("likelihood_weight")
        self.x = self.rng.normal(self.model.prior_mu, self.model.prior_sigma, size=self.n_particles)
        self.lw = self._ll([y for (y,) in self.obs])

    @property
    def log_weights(self):
        return self.lw.copy()

    def set_particle_log_weights(self, v):
        assert len(v) == self.n_particles
        self.lw = np.array(v, dtype=np.float64)

    def resample(self, P, mode="multinomial"):
        self.calls.append(("resample", P, mode))
        w = np.exp(self.lw - self.lw.max())
        anc = self.rng.choice(self.n_particles, size=P, p=w / w.sum())
        lse = self.lw.max() + np.log(w.sum())
        self.x = self.x[anc]
        self.lw = np.full(P, lse - np.log(P))
        self.n_particles = P

    def primitive_infer(self, exp):
        self.calls.append(("primitive_infer", list(exp)))
        return np.full(self.n_particles, 2.0)

    def log_likelihood_at(self, scope, block):
        return self._ll([y for (y,) in self.obs])

    def read_site(self, j):
        return self.x.copy()

    def num_sites(self):
        return 1

    def get_seed(self):
        return self.seed

    def set_seed(self, s):
        self.seed = s

    def reset_to_prior(self):
        self.x = self.rng.normal(self.model.prior_mu, self.model.prior_sigma, size=self.n_particles)

    def close(self):
        pass


def _fake_backend(model):
    from venturecxx_b200 import backend as bk

    class FakeB200(bk.B200):
        def trace_constructor(self):
            return lambda seed: bk.B200Trace(self.model, seed, trace_set=FakeBatch(self.model, 1, seed))
    return FakeB200(model)


def test_reference_engine_runs_c1_program_on_the_batch_cpu():
    """`[assume x (normal 0 1)] [observe (normal x 1) 2] (do (resample 1000) (likelihood_weight) (resample 1000))`
    (benchmarks/Makefile:6-9) through the reference Engine / Infer / inference SPs, unchanged."""
    import venturecxx_b200 as vb
    from venturecxx_b200 import backend as bk
    e = _engine(_fake_backend(vb.NormalNormalModel(0.0, 1.0, 1.0)))
    assert type(e.model).__name__ == "B200EngineTraceSet" and isinstance(e.model.traces, bk.B200Master)
    did, val = e.assume("x", ref_engine.expression("(normal 0 1)"))
    assert val["type"] == "number"
    _, incr = e.observe(ref_engine.expression("(normal x 1)"), ref_engine.value(2.0))
    assert len(incr) == 1 and incr[0] == pytest.approx(-0.5 * np.log(2 * np.pi) - 0.5 * (2.0 - val["value"]) ** 2)
    e.infer(ref_engine.program("(do (resample 1000) (likelihood_weight) (resample 1000))"))
    batch = e.model.b200
    assert batch.calls == [("resample", 1000, "multinomial"), "likelihood_weight", ("resample", 1000, "multinomial")]
    assert e.num_traces() == 1000 and len(e.model.log_weights) == 1000
    # posterior N(1, sqrt(1/2)) (test/smc/test_particle_filter.py:61-72), loosely: the stand-in is numpy
    x = batch.read_site(0)
    assert abs(x.mean() - 1.0) < 0.15 and abs(x.std() - np.sqrt(0.5)) < 0.12
    # weights after resample: all equal to logsumexp - log P (trace_set.py:206, utils.py:152-157)
    lw = np.array(e.model.log_weights)
    assert np.ptp(lw) < 1e-12
    # the Infer accessors (engine/inference.py:218-234) see the device weights
    e.infer(ref_engine.program("(do (ws <- (particle_log_weights)) (set_particle_log_weights ws) (increment_particle_log_weights ws))"))
    np.testing.assert_allclose(e.model.log_weights, 2 * lw, rtol=1e-15)
    # transition operators reach primitive_infer with the reference's argument list (dispatch.py:50-107)
    e.infer(ref_engine.program("(resimulation_mh default one 7)"))
    op = [c for c in batch.calls if isinstance(c, tuple) and c[0] == "primitive_infer"][-1][1]
    assert op[0] in ("resimulation_mh", "mh") and op[1:] == ["default", "one", 7]
    # predict of an assumed name gives one value per particle; report_value the distinguished one
    _, vals = e.predict_all(ref_engine.expression("x"))
    assert len(vals) == 1000 and vals[5]["value"] == x[5]
    assert e.report_value(did)["value"] == x[0]


def test_directives_are_checked_against_the_compiled_program_cpu():
    import venturecxx_b200 as vb
    from venturecxx_b200.exception import VentureException
    e = _engine(_fake_backend(vb.NormalNormalModel(10.0, 1.0, 1.0)))
    with pytest.raises(Exception) as err:
        e.assume("x", ref_engine.expression("(normal 0 1)"))        # the compiled prior is N(10, 1)
    assert "does not match" in str(err.value)
    e = _engine(_fake_backend(vb.NormalNormalModel(10.0, 1.0, 1.0)))
    e.assume("mu", ref_engine.expression("(normal 10 1)"))           # any name: later expressions are renamed
    e.observe(ref_engine.expression("(normal mu 1)"), ref_engine.value(14.0))
    with pytest.raises(Exception) as err:
        e.observe(ref_engine.expression("(normal mu 2)"), ref_engine.value(14.0))
    assert "observe expression" in str(err.value)
    assert issubclass(VentureException, Exception)


def test_decoders_logistic_regression_hmm_lda_cpu():
    import venturecxx_b200 as vb
    from venturecxx_b200 import backend as bk
    from venturecxx_b200.interchange import expression, boolean, number, integer
    m = vb.LogisticRegressionModel(3, 10)
    d = bk.decoder_for(m)
    for j in range(3):
        assert d.define("b%d" % j, expression(["normal", 0.0, 1.0])) == j and d.site_of_define(j) == j
    d.define("weights", expression(["array", "b0", "b1", "b2"]))
    assert d.complete()
    row, y = d.observe(expression(["log_odds_flip", ["vector_dot", "weights", ["array", 0.5, -1.0, 2.0]]]), boolean(True), 0)
    assert row.tolist() == [0.5, -1.0, 2.0] and y.tolist() == [True]
    H = __import__("helpers").golden("hmm.json")
    hm = vb.HMMModel(H["transition_matrix"], H["emission_matrix"], 1, 20)
    dh = bk.decoder_for(hm)
    for (_k, name, exp) in hm.program([]):
        dh.define(name, exp)
    assert dh.observe(expression(["emission", ["state", 1]]), number(3), 0) == (3,)
    with pytest.raises(Exception):
        dh.observe(expression(["emission", ["state", 3]]), number(3), 1)        # times in order
    lm = vb.LDAModel(4, 9, 2, 6)
    dl = bk.decoder_for(lm)
    for (_k, name, exp) in lm.program([]):
        dl.define(name, exp)
    assert dl.observe(expression(["word", 1, 0]), integer(5), 0) == (1, 5)
    assert dl.observe(expression(["word", 1, 1]), integer(2), 1) == (1, 2)
    assert dl.site_of_define(2) == 0 and dl.site_of_define(3) == 1


def test_plugin_binds_foreign_inference_sps_with_the_reference_signatures_cpu():
    """B4: ripl.bind_foreign_inference_sp (ripl.py:938-939 -> engine.py:56-59) with the SP types of
    backend/lite/inference_sps.py:234-241, 536-551, 602-606, 678-694; a Lite-backed engine then runs
    `(b200_resimulation_mh default one 5)` -- and, with override_builtins, `(resimulation_mh default one 5)` -- on the batch."""
    import venturecxx_b200 as vb
    from venturecxx_b200 import plugin
    e = ref_engine.make_engine(seed=3)                     # the reference engine over LITE

    class Ripl(object):                                    # the three ripl methods a plugin uses (ripl.py:934-947)
        def bind_foreign_inference_sp(self, name, sp):
            e.bind_foreign_inference_sp(name, sp)

        def bind_callback(self, name, cb):
            e.bind_callback(name, cb)
    batch = FakeBatch(vb.NormalNormalModel(0.0, 1.0, 1.0), 50, 7)
    p = plugin.__venture_start__(Ripl(), trace_set=batch, override_builtins=False)
    assert p.trace_set is batch and "b200_resimulation_mh" in e.inference_sps and "b200_resample" in e.inference_sps
    e.infer(ref_engine.program("(do (b200_resample 80) (b200_likelihood_weight) (b200_resimulation_mh default one 5))"))
    assert batch.calls[0] == ("resample", 80, "multinomial") and batch.calls[1] == "likelihood_weight"
    assert batch.calls[2] == ("primitive_infer", ["resimulation_mh", "default", "one", 5])
    assert batch.n_particles == 80 and e.num_traces() == 1      # the Lite trace set of the engine is untouched
    _, ans = e.infer(ref_engine.program("(b200_subsampled_mh default one 16 3 0.01 false 0 false 2)"))
    assert batch.calls[-1] == ("primitive_infer", ["subsampled_mh", "default", "one", 16, 3, 0.01, False, 0, False, 2])
    e.infer(ref_engine.program("(do (b200_subsampled_mh_check_applicability default one 1) "
                               "(b200_subsampled_mh_make_consistent default one false 0 false 1))"))
    assert batch.calls[-2] == ("primitive_infer", ["subsampled_mh_check_applicability", "default", "one", 1])
    assert batch.calls[-1] == ("primitive_infer", ["subsampled_mh_make_consistent", "default", "one", False, 0, False, 1])
    e.infer(ref_engine.program("(do (b200_slice default 0 0.5 10 3) (b200_pgibbs (quote seq) ordered 8 2))"))
    assert batch.calls[-2] == ("primitive_infer", ["slice", "default", 0, 0.5, 10, 3])
    assert batch.calls[-1] == ("primitive_infer", ["pgibbs", "seq", "ordered", 8, 2])
    # the reference's own names, overridden
    e2 = ref_engine.make_engine(seed=4)

    class Ripl2(Ripl):
        def bind_foreign_inference_sp(self, name, sp):
            e2.bind_foreign_inference_sp(name, sp)

        def bind_callback(self, name, cb):
            e2.bind_callback(name, cb)
    batch2 = FakeBatch(vb.NormalNormalModel(0.0, 1.0, 1.0), 10, 8)
    plugin.__venture_start__(Ripl2(), trace_set=batch2, override_builtins=True)
    e2.infer(ref_engine.program("(do (resample 20) (likelihood_weight) (resimulation_mh default one 3))"))
    assert batch2.calls == [("resample", 20, "multinomial"), "likelihood_weight",
                            ("primitive_infer", ["resimulation_mh", "default", "one", 3])]


# ---- GPU: the real batch -----------------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_reference_engine_on_the_device_c1_posterior():
    import venturecxx_b200 as vb
    from venturecxx_b200 import backend as bk
    from stats_policy import report_known_gaussian, statistical_test

    @statistical_test
    def run(seed):
        e = _engine(bk.B200(vb.NormalNormalModel(0.0, 1.0, 1.0)), seed=seed)
        e.assume("x", ref_engine.expression("(normal 0 1)"))
        e.observe(ref_engine.expression("(normal x 1)"), ref_engine.value(2.0))
        e.infer(ref_engine.program("(do (resample 4000) (likelihood_weight) (resample 4000))"))
        assert e.num_traces() == 4000
        assert np.ptp(np.array(e.model.log_weights)) < 1e-9
        # resampled particles repeat their ancestors' values; MH moves every particle, after which the KS test
        # against the posterior N(1, sqrt(1/2)) (test/smc/test_particle_filter.py:61-72) applies
        e.infer(ref_engine.program("(resimulation_mh default one 40)"))
        return report_known_gaussian(1.0, np.sqrt(0.5), e.model.b200.read_site(0)[::4])
    run()


@pytest.mark.gpu
def test_reference_engine_equals_direct_c_abi_calls_logistic_regression():
    """(do (resample P) (likelihood_weight) (resimulation_mh default one K)) through engine.infer gives the
    weights and chains of the same entry points called directly (same seed): the seam adds nothing."""
    import venturecxx_b200 as vb
    from venturecxx_b200 import backend as bk
    D, N, P, K = 4, 60, 512, 24
    rng = np.random.default_rng(5)
    X = rng.normal(size=(N, D))
    y = rng.random(N) < 1 / (1 + np.exp(-X @ rng.normal(size=D)))
    model = vb.LogisticRegressionModel(D, N)
    e = _engine(bk.B200(model), seed=33)
    for j in range(D):
        e.assume("w_%d" % j, ref_engine.expression("(normal 0 1)"))
    e.assume("w", ref_engine.expression("(array %s)" % " ".join("w_%d" % j for j in range(D))))
    for i in range(N):
        e.observe(ref_engine.expression("(log_odds_flip (vector_dot w (array %s)))" % " ".join(repr(float(v)) for v in X[i])),
                  ref_engine.value(bool(y[i])))
    seed = e.model.b200.get_seed()
    _, nodes = e.infer(ref_engine.program("(do (resample %d) (likelihood_weight) (resimulation_mh default one %d))" % (P, K)))
    lw_engine = np.array(e.model.log_weights)
    w_engine = np.stack([e.model.b200.read_site(j) for j in range(D)])
    # the same entry points, directly
    ts = vb.B200TraceSet(model, 1, seed=seed)
    for i in range(N):
        ts.observe(X[i], np.array([y[i]]))
        ts.incorporate()
    ts.resample(P, mode="multinomial")
    ts.incorporate()
    ts.likelihood_weight()
    ts.primitive_infer(["resimulation_mh", "default", "one", K])
    np.testing.assert_array_equal(lw_engine, ts.log_weights)
    np.testing.assert_array_equal(w_engine, np.stack([ts.read_site(j) for j in range(D)]))
    assert len(lw_engine) == P and np.all(np.isfinite(lw_engine)) and np.ptp(w_engine[0]) > 0
    # the Infer accessors and log_likelihood_at go through the Master
    _, ll = e.infer(ref_engine.program("(log_likelihood_at default all)"))
    assert len(ll["value"] if isinstance(ll, dict) else ll) == P
    # particle 0 comes back as a REAL Lite trace holding the device state (Master.retrieve -> interchange)
    lite = e.getDistinguishedTrace()
    assert lite.trace.__class__.__module__.startswith("venture.lite")
    assert lite.report_value(1)["value"] == w_engine[0, 0]


@pytest.mark.gpu
def test_reference_engine_hmm_particle_filter_program():
    """examples/hmm.vnt's particle filter, step by step through the reference engine: observe (auto-incorporate),
    then (resample P); filtered marginals within TVD of the exact table (hmm_plugin.py:175-205, :225-232)."""
    import venturecxx_b200 as vb
    from venturecxx_b200 import backend as bk
    from helpers import golden
    H = golden("hmm.json")
    model = vb.HMMModel(H["transition_matrix"], H["emission_matrix"], 1, 20)
    e = _engine(bk.B200(model, resample_mode="systematic"), seed=5)
    for (_k, name, exp) in model.program([]):
        e.assume(name, exp)
    P = 1 << 16
    e.infer(ref_engine.program("(resample %d)" % P))
    for t, d in enumerate(H["data"]):
        e.observe(ref_engine.expression("(emission (state %d))" % (t + 1)), ref_engine.value(int(d)))
        e.infer(ref_engine.program("(resample %d)" % P))
        s = e.model.b200.read_site(t).astype(np.int64)
        m = np.bincount(s, minlength=6)[1:6] / float(P)
        assert 0.5 * np.abs(m - np.array(H["filtered_marginals"][t])).sum() < 0.02, t


@pytest.mark.gpu
def test_reference_engine_compiles_the_program_from_its_directives():
    """B200() without a hand-described model: the directives the engine evaluates are compiled (N2).  A linear
    regression -- a model with no .cu file -- through the reference Engine, against the same compiled model
    driven directly, and its posterior mean against the conjugate answer for the slope."""
    import venturecxx_b200 as vb
    from venturecxx_b200 import backend as bk
    from venturecxx_b200.compiler import CompiledModel
    rng = np.random.default_rng(2)
    xs = np.round(rng.normal(size=12), 3)
    ys = np.round(1.2 * xs + 0.1 * rng.normal(size=12), 3)
    e = _engine(bk.B200(), seed=9)
    assert e.num_traces() == 1
    e.assume("a", ref_engine.expression("(normal 0 2)"))
    e.infer(ref_engine.program("(resample 256)"))
    for x, y in zip(xs, ys):
        e.observe(ref_engine.expression("(normal (* a %r) 0.5)" % (float(x),)), ref_engine.value(float(y)))
    assert e.num_traces() == 256
    ts = e.model.b200
    assert ts.num_sites() == 1 and ts.model.n_observations == 12
    # weights accumulated observation by observation = log likelihood of all of them + the evened-out start
    ll = ts.log_likelihood_at("default", "all")
    lw = np.array(e.model.log_weights)
    np.testing.assert_allclose(lw - lw[0], ll - ll[0], rtol=1e-9, atol=1e-9)
    _, nodes = e.infer(ref_engine.program("(resimulation_mh default one 200)"))
    a = ts.read_site(0)
    prec = 1 / 4.0 + (xs ** 2).sum() / 0.25
    mean = (xs * ys).sum() / 0.25 / prec
    assert abs(a.mean() - mean) < 5 * np.sqrt(1 / prec / 256) + 0.01
    # a later assume extends the program: old sites keep their values, the new one is drawn from its prior
    e.assume("b", ref_engine.expression("(normal a 1)"))
    ts2 = e.model.b200
    assert ts2.num_sites() == 2
    np.testing.assert_array_equal(ts2.read_site(0), a)
    assert abs(np.mean(ts2.read_site(1) - a)) < 0.3 and 0.7 < np.std(ts2.read_site(1) - a) < 1.3
    # outside the subset: the error names the directive
    with pytest.raises(Exception) as err:
        e.assume("f", ref_engine.expression("(lambda (x) x)"))
    assert "outside" in str(err.value) or "subset" in str(err.value)
    assert isinstance(ts2.model, CompiledModel) and vb is not None


@pytest.mark.gpu
def test_reference_engine_runs_subsampled_mh_on_a_compiled_program():
    """The model and the command of test/inference_language/test_subsampledmh.py:64-99 through the reference's
    Engine: 100 `(observe (normal mu 1.0) x_i)` directives are compiled, `(subsampled_mh default one 2 3 0.01 false
    0 false 50)` runs on the device, and the engine gets back what Lite returns (the node count of the scaffold cut
    at the global border: 1 -- mu is its own border)."""
    from venturecxx_b200 import backend as bk
    from stats_policy import report_known_continuous, statistical_test
    import scipy.stats as st

    @statistical_test
    def check(seed):
        x = np.random.RandomState(3).normal(10, 1, 100)
        e = _engine(bk.B200(), seed=seed)
        e.assume("mu", ref_engine.expression("(normal 10.0 3.0)"))
        e.infer(ref_engine.program("(resample 128)"))
        for v in x:
            e.observe(ref_engine.expression("(normal mu 1.0)"), ref_engine.value(float(v)))
        nodes = None
        for _ in range(12):
            _, nodes = e.infer(ref_engine.program("(subsampled_mh default one 2 3 0.01 false 0 false 50)"))
        assert nodes is not None
        mean = (9.0 * x.sum() + 10.0) / (9.0 * 100 + 1.0)
        std = np.sqrt(9.0 / (9.0 * 100 + 1.0))
        return report_known_continuous(st.norm(loc=mean, scale=std).cdf, e.model.b200.read_site(0), "N(%f,%f^2)" % (mean, std))
    check()


@pytest.mark.gpu
def test_reference_engine_compiles_tagged_scopes():
    """`(tag (quote wts) j ...)` directives and `(resimulation_mh (quote wts) j K)` / `each` commands through the reference's
    Engine: only the tagged block moves"""
    from venturecxx_b200 import backend as bk
    e = _engine(bk.B200(), seed=5)
    e.assume("a", ref_engine.expression("(tag (quote wts) 0 (normal 0 1))"))
    e.assume("b", ref_engine.expression("(tag (quote wts) 1 (normal a 1))"))
    e.assume("c", ref_engine.expression("(normal b 1)"))
    e.infer(ref_engine.program("(resample 64)"))
    e.observe(ref_engine.expression("(normal c 0.5)"), ref_engine.value(1.5))
    ts = e.model.b200
    assert ts.model.scopes == {"default": 0, "wts": 1} and ts.num_sites() == 3
    before = np.stack([ts.read_site(j) for j in range(3)])
    _, nodes = e.infer(ref_engine.program("(resimulation_mh (quote wts) 1 30)"))
    after = np.stack([ts.read_site(j) for j in range(3)])
    np.testing.assert_array_equal(after[0], before[0]); np.testing.assert_array_equal(after[2], before[2])
    assert (after[1] != before[1]).any()
    e.infer(ref_engine.program("(resimulation_mh (quote wts) each 5)"))
    final = np.stack([ts.read_site(j) for j in range(3)])
    np.testing.assert_array_equal(final[2], before[2])
    assert (final[0] != before[0]).any()


@pytest.mark.gpu
def test_reference_engine_runs_the_logistic_regression_program_text():
    """INTEGRATION.md's example: tagged weights, `(array ...)`, `vector_dot`, boolean observations, `each` over the tagged
    scope and `subsampled_mh` -- the reference Engine on the compiled program; the posterior mean of the weight with the
    most informative column has the data's sign"""
    from venturecxx_b200 import backend as bk
    rng = np.random.default_rng(4)
    X = np.round(rng.normal(size=(80, 2)), 3)
    y = (X @ np.array([2.0, -1.0]) + 0.3 * rng.normal(size=80)) > 0
    e = _engine(bk.B200(), seed=6)
    e.assume("w_0", ref_engine.expression("(tag (quote wts) 0 (normal 0 1))"))
    e.assume("w_1", ref_engine.expression("(tag (quote wts) 1 (normal 0 1))"))
    e.assume("w", ref_engine.expression("(array w_0 w_1)"))
    e.infer(ref_engine.program("(resample 256)"))
    for x, yi in zip(X, y):
        e.observe(ref_engine.expression("(log_odds_flip (vector_dot w (array %r %r)))" % (float(x[0]), float(x[1]))),
                  ref_engine.value(bool(yi)))
    ts = e.model.b200
    assert ts.num_sites() == 2 and ts.model.n_observations == 80 and ts.model.prog.subsampled[0]["kind"] == 1
    e.infer(ref_engine.program("(do (resimulation_mh (quote wts) each 40) (subsampled_mh default one 16 3 0.01 false 0 false 20))"))
    w0, w1 = ts.read_site(0), ts.read_site(1)
    assert w0.mean() > 0.5 and w1.mean() < 0.0
