"""N2, GPU: compiled programs (venturecxx_b200/compiler.py -> csrc/k_program.cu) against the reference.

  * the three programs of tests/golden/program_chains.json (produced by running Lite itself): simulated values,
    likelihood weights, increments, MH chains, node counts, log likelihood / joint at 1e-12;
  * a compiled normal-normal program reproduces the hand-written kernels (k_normal.cu) bit for bit and the
    Lite fixture of the C1 model;
  * one warp per particle (many observations) equals one thread per particle;
  * posterior of a coin by KS under the reference's policy; resampling; restoring device particles into Lite.
"""
import os
import sys

import numpy as np
import pytest
import scipy.stats as st

import venturecxx_b200 as vb
from helpers import golden, unj
from venturecxx_b200.compiler import CompiledModel

pytestmark = pytest.mark.gpu
RTOL = 1e-12


def sites_of(ts):
    return np.stack([ts.read_site(j) for j in range(ts.num_sites())])


@pytest.mark.parametrize("name", ["regression", "coin", "hierarchy", "mixture3"])
def test_compiled_program_reproduces_lite_on_the_device(name):
    case = golden("program_chains.json")[name]
    prog = [tuple(d) for d in case["program"]]
    assumes = [d for d in prog if d[0] == "assume"]
    observes = [d for d in prog if d[0] == "observe"]
    n_first, P, K, seed = case["n_first"], case["P"], case["K"], case["seed"]
    recs = case["particles"]
    ts = vb.B200TraceSet(CompiledModel(assumes + observes[:n_first]), P, seed=seed)
    ts.incorporate()
    ts.rng_step = case["steps"]["likelihood_weight"]
    ts.likelihood_weight()
    if name == "mixture3":
        # categorical sites, `simplex`, `lookup`.  Lite regenerates the choices in the order of its scaffold's border, here
        # not the order of the assumes (same law, other stream positions): start from Lite's prior values
        want = np.array([[unj(v) for v in r["values_lw"]] for r in recs]).T
        for j in range(ts.num_sites()):
            ts.write_site(j, want[j])
        ts.log_weights = ts.log_likelihood_at("default", "all")
    np.testing.assert_allclose(ts.log_weights, [unj(r["lw"]) for r in recs], rtol=RTOL)
    s0 = sites_of(ts)
    np.testing.assert_allclose(s0.T, [[unj(v) for v in r["values_lw"]] for r in recs], rtol=RTOL)
    # the remaining observations arrive (a longer program, same sites)
    ts2 = vb.B200TraceSet(CompiledModel(prog), P, seed=seed)
    for j in range(ts2.num_sites()):
        ts2.write_site(j, s0[j])
    incr = ts2.log_likelihood_at("default", "all") - ts.log_likelihood_at("default", "all")
    np.testing.assert_allclose(incr, [unj(r["incr"]) for r in recs], rtol=1e-10)
    inc2 = ts2.incorporate()
    np.testing.assert_allclose(inc2, ts2.log_likelihood_at("default", "all"), rtol=RTOL)
    ts2.rng_step = case["steps"]["mh"]
    nodes, acc = ts2.resimulation_mh("default", "one", K, return_accepts=True)
    np.testing.assert_allclose(sites_of(ts2).T, [[unj(v) for v in r["values_mh"]] for r in recs], rtol=RTOL)
    np.testing.assert_allclose(nodes, [np.mean(r["nodes"]) for r in recs], rtol=1e-14)
    assert acc.sum() > 0 and ts2.rng_step == case["steps"]["mh"] + K
    np.testing.assert_allclose(ts2.log_likelihood_at("default", "all"), [unj(r["ll_at"]) for r in recs], rtol=RTOL)
    np.testing.assert_allclose(ts2.log_joint_at("default", "all"), [unj(r["lj_at"]) for r in recs], rtol=RTOL)


@pytest.mark.parametrize("name", ["regression_slice_doubling", "chain_slice_doubling"])
def test_compiled_slice_doubling_reproduces_lite_on_the_device(name):
    """`(slice_doubling default one w p K)` on the device against Lite's DoublingSliceOperator run on the same uniforms
    (infer/slice_sample.py:148-196: doubling of the interval, Neal's acceptance test)"""
    case = golden("program_chains.json")[name]
    recs = case["particles"]
    ts = vb.B200TraceSet(CompiledModel([tuple(d) for d in case["program"]]), case["P"], seed=case["seed"])
    ts.incorporate()
    start = np.array([[unj(v) for v in r["values_start"]] for r in recs]).T
    for j in range(ts.num_sites()):
        ts.write_site(j, start[j])
    ts.rng_step = case["steps"]["slice"]
    nodes, moved = ts.slice_doubling("default", "one", case["w"], case["m"], case["K"], return_moves=True)
    np.testing.assert_allclose(sites_of(ts).T, [[unj(v) for v in r["values_slice"]] for r in recs], rtol=1e-11)
    np.testing.assert_allclose(nodes, [np.mean(r["nodes"]) for r in recs], rtol=1e-14)
    np.testing.assert_allclose(ts.log_joint_at("default", "all"), [unj(r["lj_at"]) for r in recs], rtol=1e-11)
    assert moved.min() > 0
    out = ts.primitive_infer(["slice_doubling", "default", "each", case["w"], case["m"], 2])
    assert out.shape == (case["P"],)


@pytest.mark.parametrize("name", ["regression_slice", "chain_slice"])
def test_compiled_slice_reproduces_lite_on_the_device(name):
    """N4 (slice): `(slice default one w m K)` on the device against Lite's StepOutSliceOperator run on the same
    uniforms (infer/slice_sample.py:39-139; fixture by tests/golden/make_program_chains.py)"""
    case = golden("program_chains.json")[name]
    recs = case["particles"]
    ts = vb.B200TraceSet(CompiledModel([tuple(d) for d in case["program"]]), case["P"], seed=case["seed"])
    ts.incorporate()
    start = np.array([[unj(v) for v in r["values_start"]] for r in recs]).T
    for j in range(ts.num_sites()):
        ts.write_site(j, start[j])
    ts.rng_step = case["steps"]["slice"]
    nodes, moved = ts.slice("default", "one", case["w"], case["m"], case["K"], return_moves=True)
    np.testing.assert_allclose(sites_of(ts).T, [[unj(v) for v in r["values_slice"]] for r in recs], rtol=1e-11)
    np.testing.assert_allclose(nodes, [np.mean(r["nodes"]) for r in recs], rtol=1e-14)
    np.testing.assert_allclose(ts.log_joint_at("default", "all"), [unj(r["lj_at"]) for r in recs], rtol=1e-11)
    assert moved.min() > 0 and ts.rng_step == case["steps"]["slice"] + case["K"]
    # through primitive_infer, as the engine hands it down (dispatch.py:190-194)
    out = ts.primitive_infer(["slice", "default", "each", case["w"], case["m"], 2])
    assert out.shape == (case["P"],)
    hand = vb.B200TraceSet(vb.NormalNormalModel(), 4)
    with pytest.raises(vb.VentureException):
        hand.primitive_infer(["slice", "default", "one", 1.0, 5, 1])


def test_slice_posterior_of_a_normal_mean():
    """slice sampling leaves the posterior invariant: prior N(10, 1), observation 14 with noise 1 => N(12, sqrt(1/2))
    (test/inference_quality/micro/test_basic_stats.py:65-76 runs the same model under every operator)"""
    from stats_policy import report_known_gaussian, statistical_test

    @statistical_test
    def run(seed):
        ts = vb.B200TraceSet(CompiledModel([("assume", "x", ["normal", 10.0, 1.0]), ("observe", ["normal", "x", 1.0], 14.0)]), 5000, seed=seed)
        ts.incorporate()
        ts.slice("default", "one", 1.0, 20, 30)
        return report_known_gaussian(12.0, np.sqrt(0.5), ts.read_site(0))
    run()


def test_compiled_normal_normal_equals_the_hand_written_kernels():
    mu0, s0, s, ys = 0.5, 2.0, 1.5, [2.0, 1.0, -0.5]
    P = 3001
    hand = vb.B200TraceSet(vb.NormalNormalModel(mu0, s0, s), P, seed=11)
    hand.observe(ys)
    gen = vb.B200TraceSet(CompiledModel([("assume", "x", ["normal", mu0, s0])] + [("observe", ["normal", "x", s], y) for y in ys]), P, seed=11)
    np.testing.assert_array_equal(gen.read_site(0), hand.read_site(0))          # the prior draw of `assume`
    np.testing.assert_array_equal(gen.incorporate(), hand.incorporate())
    hand.likelihood_weight(); gen.likelihood_weight()
    np.testing.assert_array_equal(gen.log_weights, hand.log_weights)
    n1, a1 = hand.resimulation_mh("default", "one", 50, return_accepts=True)
    n2, a2 = gen.resimulation_mh("default", "one", 50, return_accepts=True)
    np.testing.assert_array_equal(gen.read_site(0), hand.read_site(0))
    np.testing.assert_array_equal(a1, a2)
    np.testing.assert_array_equal(n1, n2)                                        # 1 + #observations (scaffold.py:67-68)
    np.testing.assert_array_equal(gen.log_likelihood_at("default", "all"), hand.log_likelihood_at("default", "all"))
    np.testing.assert_array_equal(gen.log_joint_at("default", "all"), hand.log_joint_at("default", "all"))
    anc = gen.resample(P, mode="systematic", return_ancestors=True)
    np.testing.assert_array_equal(anc, hand.resample(P, mode="systematic", return_ancestors=True))
    np.testing.assert_array_equal(gen.read_site(0), hand.read_site(0))
    # and the C1 chain Lite itself produced (tests/golden/lite_chains.json)
    c = golden("lite_chains.json")["c1"]
    g = vb.B200TraceSet(CompiledModel([("assume", "x", ["normal", c["mu0"], c["s0"]])] +
                                      [("observe", ["normal", "x", c["s"]], y) for y in c["y"]]), c["P"], seed=c["seed"])
    g.incorporate()
    g.rng_step = c["steps"]["mh"]
    # (the fixture runs likelihood_weight on two observations first; start the chain from its state instead)
    hand2 = vb.B200TraceSet(vb.NormalNormalModel(c["mu0"], c["s0"], c["s"]), c["P"], seed=c["seed"])
    hand2.observe(c["y"][:2]); hand2.incorporate()
    hand2.rng_step = c["steps"]["likelihood_weight"]
    hand2.likelihood_weight()
    g.write_site(0, hand2.read_site(0))
    g.resimulation_mh("default", "one", c["K"])
    np.testing.assert_allclose(g.read_site(0), [unj(r["x_mh"]) for r in c["particles"]], rtol=RTOL)


def test_one_warp_per_particle_equals_the_python_interpreter():
    """>= 64 observations in a group: lanes take rows in turn and the warp adds their sums (another summation
    order than Lite's, so 1e-10); chain decisions equal the sequential interpreter's"""
    from oracle.program_vm import VM
    from venturecxx_b200 import compiler
    rng = np.random.default_rng(3)
    xs = rng.normal(size=150)
    ys = 0.8 * xs + 0.3 + 0.5 * rng.normal(size=150)
    prog = [("assume", "a", ["normal", 0.0, 2.0]), ("assume", "b", ["normal", 0.0, 2.0]), ("assume", "s", ["gamma", 2.0, 2.0])] + \
           [("observe", ["normal", ["add", ["mul", "a", float(x)], "b"], "s"], float(y)) for x, y in zip(xs, ys)]
    m = CompiledModel(prog)
    P, K = 6, 40
    ts = vb.B200TraceSet(m, P, seed=5)
    inc = ts.incorporate()
    vm = VM(m.prog.blob(), compiler.OP, P, 5)
    vm.reset_to_prior()
    np.testing.assert_allclose(sites_of(ts), vm.sites, rtol=RTOL)
    np.testing.assert_allclose(inc, vm.log_likelihood(), rtol=1e-10)
    step = ts.rng_step
    nodes, acc = ts.resimulation_mh("default", "one", K, return_accepts=True)
    vm.step = step
    acc_ref, nodes_ref = vm.mh(K)
    np.testing.assert_array_equal(acc, acc_ref)
    np.testing.assert_allclose(sites_of(ts), vm.sites, rtol=RTOL)
    np.testing.assert_allclose(nodes, nodes_ref, rtol=1e-14)
    np.testing.assert_allclose(ts.log_joint_at("default", "all"), vm.log_likelihood(joint=True), rtol=1e-10)
    ts.primitive_infer(["resimulation_mh", "default", "each", 3])
    vm.mh(3 * 3, block=-3)
    np.testing.assert_allclose(sites_of(ts), vm.sites, rtol=RTOL)


def test_coin_posterior_and_global_logscore():
    """a model that has no .cu file: p ~ beta(1, 1), 100 observed flips (test/conformance/test_get_global_logscore.py:
    32-50 uses the same shape); posterior Beta(1 + heads, 1 + tails) by KS under test/stats.py's policy"""
    from stats_policy import report_known_continuous, statistical_test
    flips = [True] * 62 + [False] * 38
    prog = [("assume", "p", ["beta", 1.0, 1.0])] + [("observe", ["flip", "p"], f) for f in flips]
    fair = [("assume", "x", ["normal", 0.0, 1.0])] + [("observe", ["flip", 0.5], True)] * 100
    t0 = vb.B200TraceSet(CompiledModel(fair), 4, seed=1)
    np.testing.assert_allclose(t0.incorporate(), -100 * np.log(2), rtol=1e-13)      # test_global_logscore: -100 log 2

    @statistical_test
    def run(seed):
        ts = vb.B200TraceSet(CompiledModel(prog), 4000, seed=seed)
        ts.incorporate()
        ts.resimulation_mh("default", "one", 120)
        return report_known_continuous(st.beta(63, 39).cdf, ts.read_site(0), "beta(63, 39)")
    run()


def test_compiled_particles_restore_into_lite():
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import ref_engine
    if not ref_engine.available():
        pytest.skip("oracle/_ref not built")
    ref_engine._import()
    import venture.engine.trace as tr
    import venture.lite.value as vv
    from venture.lite.trace import Trace
    case = golden("program_chains.json")["hierarchy"]
    m = CompiledModel([tuple(d) for d in case["program"]])
    ts = vb.B200TraceSet(m, 7, seed=3)
    ts.incorporate()
    ts.resimulation_mh("default", "one", 25)
    lj = ts.log_joint_at("default", "all")
    for i in (0, 3, 6):
        lite = tr.Trace.restore(lambda: Trace(1), ts.dump_particle(i), {})
        got = lite.trace.log_joint_at(vv.VentureSymbol("default"), vv.VentureSymbol("all"))
        assert abs(got - lj[i]) <= 1e-12 * abs(lj[i])


@pytest.mark.parametrize("name", ["normal_observes_ssmh", "scaled_ssmh"])
def test_compiled_subsampled_mh_reproduces_lite_on_the_device(name):
    """`(subsampled_mh default one Nbatch k0 eps false 0 false 1)` x K on the device against Lite's subsampledMixMH /
    sequentialTest (infer/subsampled_mh.py:76-171) on the same draws and the same order of local sections: every
    decision, the sections each test visited, its t statistic, the node counts, the final values"""
    case = golden("program_chains.json")[name]
    recs = case["particles"]
    ts = vb.B200TraceSet(CompiledModel([tuple(d) for d in case["program"]]), case["P"], seed=case["seed"])
    ts.rng_step = case["steps"]["prior"]
    ts.reset_to_prior()
    ts.incorporate()
    np.testing.assert_allclose(sites_of(ts).T, [[unj(v) for v in r["values_prior"]] for r in recs], rtol=RTOL)
    ts.rng_step = case["steps"]["ssmh"]
    acc, visited, tstat, nodes = [], [], [], []
    for _ in range(case["K"]):
        st_ = ts.primitive_infer(["subsampled_mh", "default", "one", case["Nbatch"], case["k0"], case["eps"], False, 0, False, 1])
        acc.append(st_[:, 0]), visited.append(st_[:, 1]), tstat.append(st_[:, 2]), nodes.append(ts.last_mean_nodes)
    acc, visited, tstat, nodes = [np.array(a).T for a in (acc, visited, tstat, nodes)]
    for p, r in enumerate(recs):
        with_border = visited[p] > 0
        np.testing.assert_array_equal(visited[p][with_border], r["n"])
        np.testing.assert_array_equal(acc[p][with_border] > 0, r["accept"])
        np.testing.assert_allclose(tstat[p][with_border], [unj(v) for v in r["tstat"]], rtol=1e-9)
        np.testing.assert_array_equal(nodes[p], r["nodes"])
    np.testing.assert_allclose(sites_of(ts).T, [[unj(v) for v in r["values"]] for r in recs], rtol=RTOL)
    assert ts.rng_step == case["steps"]["ssmh"] + case["K"]


def test_compiled_subsampled_mh_refuses_what_lite_rejects():
    bad = golden("program_chains.json")["invalid_border_ssmh"]
    ts = vb.B200TraceSet(CompiledModel([tuple(d) for d in bad["program"]]), 4, seed=1)
    ts.reset_to_prior()
    ts.incorporate()
    with pytest.raises(vb.VentureException) as e:
        ts.subsampled_mh("default", "one", 1, 1, 0.1, False, 0, False, 3)
    assert "Invalid global border" in str(e.value)
    ts.subsampled_mh("default", 1, 1, 1, 0.1, False, 0, False, 3)          # x alone has a valid border (itself)


def normal_with_observes(N, sigma, seed):
    """setupNormalWithObserves, test/inference_language/test_subsampledmh.py:87-99"""
    x = np.random.RandomState(seed).normal(10, 1, N)
    prog = [("assume", "mu", ["normal", 10.0, sigma])] + [("observe", ["normal", "mu", 1.0], float(v)) for v in x]
    sigma2 = sigma ** 2
    post_mean = (sigma2 * x.sum() + 10.0) / (sigma2 * N + 1.0)
    post_std = np.sqrt(sigma2 / (sigma2 * N + 1.0))
    return prog, post_mean, post_std


@pytest.mark.parametrize("warp", [False, True])
def test_normal_with_observes_posterior(warp):
    """The reference's own statistical pin of subsampled MH (testNormalWithObserves1,
    test/inference_language/test_subsampledmh.py:64-72): `(subsampled_mh default one 2 3 0.01 false 0 false 50)`,
    the samples of mu against N(post_mean, post_std^2) by reportKnownContinuous under the policy of test/stats.py.
    warp: 100 observations run one warp per particle (>= 64 rows), 40 one thread per particle."""
    from stats_policy import report_known_continuous, statistical_test

    @statistical_test
    def check(seed):
        prog, mean, std = normal_with_observes(100 if warp else 40, 3.0, 11)
        # collectSamples continues ONE chain, 50 transitions between samples; here 200 independent chains run the same
        # command 12 times from the prior (prior proposals are accepted about 3 % of the time: 50 transitions are no burn-in)
        ts = vb.B200TraceSet(CompiledModel(prog), 200, seed=seed)
        ts.reset_to_prior()
        ts.incorporate()
        for _ in range(12):
            ts.subsampled_mh("default", "one", 2, 3, 0.01, False, 0, False, 50)
        return report_known_continuous(st.norm(loc=mean, scale=std).cdf, ts.read_site(0), "N(%f,%f^2)" % (mean, std))
    check()


def test_compiled_subsampled_mh_with_epsilon_zero_is_exact_mh():
    """epsilon = 0: every local section is visited and the decision is the exact MH one (subsampled_mh.py:66-68):
    the chain equals resimulation_mh's on the same draws"""
    prog, _, _ = normal_with_observes(80, 3.0, 5)
    a = vb.B200TraceSet(CompiledModel(prog), 64, seed=9)
    b = vb.B200TraceSet(CompiledModel(prog), 64, seed=9)
    for ts in (a, b):
        ts.rng_step = 0
        ts.reset_to_prior()
        ts.incorporate()
        ts.rng_step = 100
    st_ = a.subsampled_mh("default", "one", 7, 2, 0.0, False, 0, False, 30)
    _, acc = b.resimulation_mh("default", "one", 30, return_accepts=True)
    np.testing.assert_array_equal(st_[:, 0], acc)
    np.testing.assert_array_equal(st_[:, 1], 80.0)
    np.testing.assert_allclose(a.read_site(0), b.read_site(0), rtol=1e-15)


def test_subsampled_mh_companion_operators():
    """`subsampled_mh_check_applicability` warns as the reference does and changes nothing (dispatch.py:153-157);
    `subsampled_mh_make_consistent` has nothing to refresh on the device and reports the global sections' node counts"""
    import warnings
    from venturecxx_b200.exception import SubsampledScaffoldNotApplicableWarning, SubsampledScaffoldNotEffectiveWarning
    case = golden("program_chains.json")
    ts = vb.B200TraceSet(CompiledModel([tuple(d) for d in case["invalid_border_ssmh"]["program"]]), 8, seed=2)
    before = sites_of(ts).copy()
    with warnings.catch_warnings(record=True) as caught:
        warnings.simplefilter("always")
        out = ts.primitive_infer(["subsampled_mh_check_applicability", "default", "one", 1])
    assert [type(w.message) for w in caught] == [SubsampledScaffoldNotApplicableWarning] and not out.any()
    np.testing.assert_array_equal(sites_of(ts), before)
    ts = vb.B200TraceSet(CompiledModel([tuple(d) for d in case["scaled_ssmh"]["program"]]), 8, seed=2)
    with warnings.catch_warnings(record=True) as caught:
        warnings.simplefilter("always")
        ts.primitive_infer(["subsampled_mh_check_applicability", "default", "one", 1])
        ts.primitive_infer(["subsampled_mh_check_applicability", "default", 0, 1])
    assert [type(w.message) for w in caught] == [SubsampledScaffoldNotEffectiveWarning]
    np.testing.assert_array_equal(ts.primitive_infer(["subsampled_mh_make_consistent", "default", "one", False, 0, False, 1]), 4.0)
    blr = vb.B200TraceSet(vb.LogisticRegressionModel(3, 16), 8, seed=2)
    with warnings.catch_warnings(record=True) as caught:
        warnings.simplefilter("always")
        blr.primitive_infer(["subsampled_mh_check_applicability", "default", "one", 1])
    assert not caught
    np.testing.assert_array_equal(blr.primitive_infer(["subsampled_mh_make_consistent", "default", 1, False, 0, False, 1]), 3.0)
    with pytest.raises(vb.VentureException):
        vb.B200TraceSet(vb.NormalNormalModel(0.0, 1.0, 1.0, 2), 4).primitive_infer(["subsampled_mh_check_applicability", "default", "one", 1])
    # mh_kernel_update (dispatch.py:145-148): the same chain as resimulation_mh on the same draws
    prog = [tuple(d) for d in case["normal_observes_ssmh"]["program"]]
    a, b = vb.B200TraceSet(CompiledModel(prog), 16, seed=4), vb.B200TraceSet(CompiledModel(prog), 16, seed=4)
    for t_ in (a, b):
        t_.incorporate()
        t_.rng_step = 50
    na = a.primitive_infer(["mh_kernel_update", "default", "one", False, 0, True, 20])
    nb = b.primitive_infer(["resimulation_mh", "default", "one", 20])
    np.testing.assert_array_equal(a.read_site(0), b.read_site(0))
    np.testing.assert_array_equal(na, nb)


def test_compiled_logistic_regression_on_the_device():
    """The C2 / C5 program TEXT (`(array ...)`, `vector_dot`, `log_odds_flip`) through the compiler and the generic kernels:
    the Lite chains of lite_chains.json c2 / c5 (the fixtures that pin the hand-written logistic-regression kernels) and the
    hand-written kernels themselves on the same seed."""
    from test_program_compiler import blr_program
    c = golden("lite_chains.json")["c2"]
    D, N, P, K, seed = c["D"], c["N"], c["P"], c["K"], c["seed"]
    recs = c["particles"]
    ts = vb.B200TraceSet(CompiledModel(blr_program(c)), P, seed=seed)
    start = np.array([r["w_lw"] for r in recs]).T
    for j in range(D):
        ts.write_site(j, start[j])
    ts.incorporate()
    ts.rng_step = c["steps"]["mh_one"]
    nodes = ts.resimulation_mh("default", "one", K)
    np.testing.assert_allclose(sites_of(ts).T, [r["w_one"] for r in recs], rtol=RTOL)
    np.testing.assert_array_equal(nodes, c["untagged_nodes_one"][0])
    np.testing.assert_allclose(ts.log_likelihood_at("default", "all"), [unj(r["ll_one"]) for r in recs], rtol=RTOL)
    # the hand-written kernels from the same start on the same Philox steps
    hand = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), P, seed=seed)
    hand.observe(np.array(c["X"]), np.array(c["y"]))
    for j in range(D):
        hand.write_site(j, start[j])
    hand.incorporate()
    hand.rng_step = c["steps"]["mh_one"]
    hand.resimulation_mh("default", "one", K)
    np.testing.assert_allclose(sites_of(ts), sites_of(hand), rtol=RTOL)
    # C5: subsampled MH, the border is the `w` array node
    c = golden("lite_chains.json")["c5"]
    recs = c["particles"]
    ts = vb.B200TraceSet(CompiledModel(blr_program(c)), c["P"], seed=c["seed"])
    ts.rng_step = c["steps"]["prior"]
    ts.reset_to_prior()
    ts.incorporate()
    np.testing.assert_allclose(sites_of(ts).T, [r["w_prior"] for r in recs], rtol=RTOL)
    ts.rng_step = c["steps"]["ssmh"]
    acc, visited = [], []
    for _ in range(c["K"]):
        st_ = ts.subsampled_mh("default", "one", c["Nbatch"], c["k0"], c["eps"], False, 0, False, 1)
        acc.append(st_[:, 0]), visited.append(st_[:, 1])
        np.testing.assert_array_equal(ts.last_mean_nodes, 3.0)
    acc, visited = np.array(acc).T, np.array(visited).T
    for p, r in enumerate(recs):
        np.testing.assert_array_equal(visited[p], r["n"])
        np.testing.assert_array_equal(acc[p] > 0, r["accept"])
    np.testing.assert_allclose(sites_of(ts).T, [r["w"] for r in recs], rtol=RTOL)


def test_tagged_scopes_on_the_device():
    """`(tag (quote wts) j (normal 0 1))`: the model lite_chains.json c2 actually ran -- node counts 32, `(resimulation_mh wts
    <block> 1)` block by block, `each` over the scope; a scope over a subset of the sites leaves the others alone"""
    from test_program_compiler import tagged_blr_program
    c = golden("lite_chains.json")["c2"]
    D, P, K, seed = c["D"], c["P"], c["K"], c["seed"]
    recs = c["particles"]
    ts = vb.B200TraceSet(CompiledModel(tagged_blr_program(c)), P, seed=seed)
    start = np.array([r["w_lw"] for r in recs]).T
    for j in range(D):
        ts.write_site(j, start[j])
    ts.incorporate()
    ts.rng_step = c["steps"]["mh_one"]
    nodes = ts.resimulation_mh("default", "one", K)
    np.testing.assert_allclose(sites_of(ts).T, [r["w_one"] for r in recs], rtol=RTOL)
    np.testing.assert_array_equal(nodes, [np.mean(r["nodes_one"]) for r in recs])
    for k in range(2 * D):
        ts.rng_step = c["steps"]["mh_each"] + k
        nodes = ts.primitive_infer(["resimulation_mh", "wts", k % D, 1])
        np.testing.assert_array_equal(nodes, [r["nodes_each"][k] for r in recs])
    np.testing.assert_allclose(sites_of(ts).T, [r["w_each"] for r in recs], rtol=RTOL)
    assert ts.numBlocksInScope("default") == D
    # `each` over the tagged scope = its blocks in order: the same chain as issuing them one by one
    a, b = [vb.B200TraceSet(CompiledModel(tagged_blr_program(c)), P, seed=seed) for _ in range(2)]
    for t_ in (a, b):
        t_.incorporate()
        t_.rng_step = 300
    na = a.primitive_infer(["resimulation_mh", "wts", "each", 2])
    for k in range(2 * D):
        b.primitive_infer(["resimulation_mh", "wts", float(k % D), 1])
    np.testing.assert_array_equal(sites_of(a), sites_of(b))
    np.testing.assert_array_equal(na, 32.0)
    sub = vb.B200TraceSet(CompiledModel([("assume", "a", ["normal", 0.0, 1.0]),
                                         ("assume", "b", ["tag", ["quote", "s"], ["quote", "x"], ["normal", "a", 1.0]]),
                                         ("assume", "c", ["tag", ["quote", "s"], 7, ["normal", "b", 1.0]]),
                                         ("observe", ["normal", "c", 1.0], 0.4)]), 64, seed=5)
    sub.incorporate()
    before = sites_of(sub).copy()
    _, acc = sub.resimulation_mh("s", "one", 60, return_accepts=True)
    sub.primitive_infer(["slice", "s", "x", 0.5, 4, 3])
    after = sites_of(sub)
    np.testing.assert_array_equal(after[0], before[0])
    assert acc.sum() > 0 and (after[1] != before[1]).any() and (after[2] != before[2]).any()
    with pytest.raises(vb.VentureException) as e:
        sub.resimulation_mh("s", 3, 1)
    assert "no block" in str(e.value)


def test_compiled_emap_reproduces_lite_on_the_device():
    """`(emap z one 1)` / `(emap z <block> 1)`: EnumerativeMAPOperator (infer/egibbs.py:117-121) on the device vs Lite"""
    case = golden("program_chains.json")["mixture_emap"]
    recs = case["particles"]
    ts = vb.B200TraceSet(CompiledModel([tuple(d) for d in case["program"]]), case["P"], seed=case["seed"])
    ts.rng_step = case["steps"]["prior"]
    ts.reset_to_prior()
    ts.incorporate()
    ts.rng_step = case["steps"]["gibbs_one"]
    nodes = [ts.primitive_infer(["emap", "z", "one", 1]) for _ in range(case["K"])]
    np.testing.assert_array_equal(np.array(nodes).T, [r["nodes_one"] for r in recs])
    np.testing.assert_array_equal(sites_of(ts).T[:, 2:], np.array([r["values_one"] for r in recs])[:, 2:])
    for k in range(2 * case["n_blocks"]):
        ts.rng_step = case["steps"]["gibbs_block"] + k
        ts.emap("z", k % case["n_blocks"], 1)
    np.testing.assert_array_equal(sites_of(ts).T[:, 2:], np.array([r["values_block"] for r in recs])[:, 2:])
    np.testing.assert_allclose(ts.log_joint_at("default", "all"), [unj(r["lj_at"]) for r in recs], rtol=RTOL)
    # a fixed point: another sweep changes nothing
    _, ch = ts.emap("z", "each", 1, return_changes=True)
    assert not ch.any()
    # gibbs_update (dispatch.py:185-189) is gibbs on the device: the same chain on the same draws
    a, b = [vb.B200TraceSet(CompiledModel([tuple(d) for d in case["program"]]), 32, seed=3) for _ in range(2)]
    for t_ in (a, b):
        t_.incorporate()
        t_.rng_step = 77
    a.primitive_infer(["gibbs_update", "z", "one", 25]), b.primitive_infer(["gibbs", "z", "one", 25])
    np.testing.assert_array_equal(sites_of(a), sites_of(b))


def test_compiled_gibbs_reproduces_lite_on_the_device():
    """`(gibbs z one 1)` x K and `(gibbs z <block> 1)` on the device against Lite's EnumerativeGibbsOperator on the same
    draws (infer/egibbs.py:42-107): a mixture with `biplex`, the three flip SPs, a flip whose probability hangs on another"""
    case = golden("program_chains.json")["mixture_gibbs"]
    recs = case["particles"]
    ts = vb.B200TraceSet(CompiledModel([tuple(d) for d in case["program"]]), case["P"], seed=case["seed"])
    ts.rng_step = case["steps"]["prior"]
    ts.reset_to_prior()
    ts.incorporate()
    np.testing.assert_allclose(sites_of(ts).T, [r["values_prior"] for r in recs], rtol=RTOL)
    ts.rng_step = case["steps"]["gibbs_one"]
    nodes = [ts.primitive_infer(["gibbs", "z", "one", 1]) for _ in range(case["K"])]
    np.testing.assert_array_equal(np.array(nodes).T, [r["nodes_one"] for r in recs])
    np.testing.assert_array_equal(sites_of(ts).T[:, 2:], np.array([r["values_one"] for r in recs])[:, 2:])     # the assignments
    np.testing.assert_allclose(sites_of(ts).T[:, :2], np.array([r["values_one"] for r in recs])[:, :2], rtol=RTOL)
    nodes = []
    for k in range(2 * case["n_blocks"]):
        ts.rng_step = case["steps"]["gibbs_block"] + k
        nodes.append(ts.primitive_infer(["gibbs", "z", k % case["n_blocks"], 1]))
    np.testing.assert_array_equal(np.array(nodes).T, [r["nodes_block"] for r in recs])
    np.testing.assert_array_equal(sites_of(ts).T[:, 2:], np.array([r["values_block"] for r in recs])[:, 2:])
    np.testing.assert_allclose(ts.log_joint_at("default", "all"), [unj(r["lj_at"]) for r in recs], rtol=RTOL)
    # continuous sites cannot be enumerated (PSP.enumerateValues raises in the reference)
    with pytest.raises(vb.VentureException) as e:
        ts.primitive_infer(["gibbs", "default", "one", 1])
    assert "not boolean" in str(e.value)
    # the posterior of one assignment by enumeration: z4 given the means (its observation is 0.2, flip 0.5 prior)
    big = vb.B200TraceSet(CompiledModel([tuple(d) for d in case["program"]]), 4000, seed=8)
    big.incorporate()
    big.write_site(0, np.full(4000, -1.0)); big.write_site(1, np.full(4000, 2.0))
    big.primitive_infer(["gibbs", "z", 4, 3])
    l1, l0 = st.norm(2.0, 0.7).pdf(0.2), st.norm(-1.0, 0.7).pdf(0.2)
    want = l1 / (l1 + l0)
    assert abs(big.read_site(6).mean() - want) < 4 * np.sqrt(want * (1 - want) / 4000) + 1e-3


def test_compiled_gibbs_over_categorical_sites_on_the_device():
    """`(gibbs z one 1)` x K and per block on categorical assignments (weights literal and hanging on a beta site, explicit
    objects, `lookup` into the array of means) against Lite's EnumerativeGibbsOperator on the same draws"""
    case = golden("program_chains.json")["mixture3_gibbs"]
    recs = case["particles"]
    ts = vb.B200TraceSet(CompiledModel([tuple(d) for d in case["program"]]), case["P"], seed=case["seed"])
    ts.rng_step = case["steps"]["prior"]
    ts.reset_to_prior()
    ts.incorporate()
    np.testing.assert_allclose(sites_of(ts).T, [r["values_prior"] for r in recs], rtol=RTOL)
    zs = [j for j, n in enumerate(recs[0]["sites"]) if n.startswith("z")]
    ts.rng_step = case["steps"]["gibbs_one"]
    nodes = [ts.primitive_infer(["gibbs", "z", "one", 1]) for _ in range(case["K"])]
    np.testing.assert_array_equal(np.array(nodes).T, [r["nodes_one"] for r in recs])
    np.testing.assert_array_equal(sites_of(ts).T[:, zs], np.array([r["values_one"] for r in recs])[:, zs])
    for k in range(2 * case["n_blocks"]):
        ts.rng_step = case["steps"]["gibbs_block"] + k
        ts.gibbs("z", k % case["n_blocks"], 1)
    np.testing.assert_array_equal(sites_of(ts).T[:, zs], np.array([r["values_block"] for r in recs])[:, zs])
    np.testing.assert_allclose(ts.log_joint_at("default", "all"), [unj(r["lj_at"]) for r in recs], rtol=RTOL)
