"""N2, CPU: the model compiler (venturecxx_b200/compiler.py) against the reference's Lite backend.

tests/golden/program_chains.json was produced by running Lite itself on three programs (linear regression,
a beta-Bernoulli coin, a hierarchy with latent children, deterministic assumes and six SP kinds) with the
oracle's draws behind Lite's RNGs (tests/golden/make_program_chains.py).  Here the SAME programs are compiled
and the compiled code is executed by the pure-Python interpreter oracle/program_vm.py: simulated values,
likelihood weights, incorporate increments, every MH decision, node counts and log likelihoods must equal
Lite's.  tests/test_gpu_program.py then holds the kernels to the same fixture.
"""
import numpy as np
import pytest

import oracle
from helpers import golden, unj
from oracle.program_vm import VM
from venturecxx_b200 import compiler
from venturecxx_b200.compiler import CompiledModel
from venturecxx_b200.exception import VentureException

RTOL = 1e-12


def split(case):
    prog = [tuple(d) for d in case["program"]]
    assumes = [d for d in prog if d[0] == "assume"]
    observes = [d for d in prog if d[0] == "observe"]
    return assumes, observes


@pytest.mark.parametrize("name", ["regression", "coin", "hierarchy", "mixture3"])
def test_compiled_program_reproduces_lite(name):
    case = golden("program_chains.json")[name]
    assumes, observes = split(case)
    n_first, P, K, seed = case["n_first"], case["P"], case["K"], case["seed"]
    first = CompiledModel(assumes + observes[:n_first])
    full = CompiledModel(assumes + observes)
    assert first.site_names == case["particles"][0]["sites"] == full.site_names
    vm = VM(first.prog.blob(), compiler.OP, P, seed)
    vm.step = case["steps"]["likelihood_weight"]
    vm.likelihood_weight()
    recs = case["particles"]
    if name == "mixture3":
        # Lite regenerates the choices in the order of its scaffold's border, here not the order of the assumes: same law,
        # other stream positions.  Start from Lite's prior values; the weights are then their log likelihood.
        vm.sites[:] = np.array([[unj(v) for v in r["values_lw"]] for r in recs]).T
        vm.lw[:] = vm.log_likelihood()
    np.testing.assert_allclose(vm.lw, [unj(r["lw"]) for r in recs], rtol=RTOL)
    np.testing.assert_allclose(vm.sites.T, [[unj(v) for v in r["values_lw"]] for r in recs], rtol=RTOL)
    # the remaining observations arrive: the increment is their log density (lite/trace.py:422-448)
    vm2 = VM(full.prog.blob(), compiler.OP, P, seed)
    vm2.sites[:] = vm.sites
    incr = vm2.log_likelihood() - vm.log_likelihood()
    np.testing.assert_allclose(incr, [unj(r["incr"]) for r in recs], rtol=1e-10)
    vm2.step = case["steps"]["mh"]
    acc, nodes = vm2.mh(K)
    np.testing.assert_allclose(vm2.sites.T, [[unj(v) for v in r["values_mh"]] for r in recs], rtol=RTOL)
    np.testing.assert_allclose(nodes, [np.mean(r["nodes"]) for r in recs], rtol=1e-14)
    assert acc.sum() > 0
    np.testing.assert_allclose(vm2.log_likelihood(), [unj(r["ll_at"]) for r in recs], rtol=RTOL)
    np.testing.assert_allclose(vm2.log_likelihood(joint=True), [unj(r["lj_at"]) for r in recs], rtol=RTOL)


@pytest.mark.parametrize("name", ["regression_slice", "chain_slice", "regression_slice_doubling", "chain_slice_doubling"])
def test_compiled_slice_reproduces_lite(name):
    """(slice default one w m 1) / (slice_doubling default one w p 1) x K on Lite (infer/slice_sample.py) vs the compiled
    program, same uniforms"""
    case = golden("program_chains.json")[name]
    m = CompiledModel([tuple(d) for d in case["program"]])
    recs = case["particles"]
    vm = VM(m.prog.blob(), compiler.OP, case["P"], case["seed"])
    vm.sites[:] = np.array([[unj(v) for v in r["values_start"]] for r in recs]).T
    vm.step = case["steps"]["slice"]
    nodes = vm.slice(case["K"], case["w"], case["m"], doubling=name.endswith("doubling"))
    np.testing.assert_allclose(vm.sites.T, [[unj(v) for v in r["values_slice"]] for r in recs], rtol=1e-11)
    np.testing.assert_allclose(nodes, [np.mean(r["nodes"]) for r in recs], rtol=1e-14)
    np.testing.assert_allclose(vm.log_likelihood(joint=True), [unj(r["lj_at"]) for r in recs], rtol=1e-11)


@pytest.mark.parametrize("name", ["normal_observes_ssmh", "scaled_ssmh"])
def test_compiled_subsampled_mh_reproduces_lite(name):
    """(subsampled_mh default one Nbatch k0 eps false 0 false 1) x K on Lite (infer/subsampled_mh.py:76-171, 326-455) vs the
    compiled program: the global border and the order of the local sections, every sequential test's decision, the
    number of sections it visited and its t statistic, the node counts of the cut scaffold"""
    case = golden("program_chains.json")[name]
    m = CompiledModel([tuple(d) for d in case["program"]])
    recs = case["particles"]
    vm = VM(m.prog.blob(), compiler.OP, case["P"], case["seed"])
    vm.step = case["steps"]["prior"]
    vm.reset_to_prior()
    np.testing.assert_allclose(vm.sites.T, [[unj(v) for v in r["values_prior"]] for r in recs], rtol=RTOL)
    vm.step = case["steps"]["ssmh"]
    accs, visited, nodes = [], [], []
    for _ in range(case["K"]):
        a, n, _, nd = vm.subsampled_mh(1, case["Nbatch"], case["k0"], case["eps"])
        accs.append(a), visited.append(n), nodes.append(nd)
    accs, visited, nodes = np.array(accs).T, np.array(visited).T, np.array(nodes).T
    for p, r in enumerate(recs):
        # sequentialTest ran only in the transitions whose site has a border: those report its n, the others 0
        with_border = visited[p] > 0
        assert with_border.sum() == len(r["n"])
        np.testing.assert_array_equal(visited[p][with_border], r["n"])
        np.testing.assert_array_equal(accs[p][with_border] > 0, r["accept"])
        np.testing.assert_array_equal(nodes[p], r["nodes"])
    np.testing.assert_allclose(vm.sites.T, [[unj(v) for v in r["values"]] for r in recs], rtol=RTOL)


def test_global_border_rules():
    """findGlobalBorder (infer/subsampled_mh.py:326-368) on the compiler's graph"""
    case = golden("program_chains.json")
    ss = CompiledModel([tuple(d) for d in case["scaled_ssmh"]["program"]]).prog.subsampled
    # mu: border = z (the mul), local sections = the latent child x, then the 24 observations; s: its own border;
    # lo and x: one consumer / none -> regular MH
    assert [s["kind"] for s in ss] == [1, 1, 0, 0]
    assert ss[0]["roots"] == [(-1, 3)] + [(1, r) for r in range(24)] and ss[0]["nodes"] == 3.0
    assert ss[1]["roots"] == [(1, r) for r in range(24)] and ss[1]["nodes"] == 1.0
    # Lite raises SubsampledScaffoldError("Invalid global border.") when the walk passes a random choice (:177-180)
    bad = case["invalid_border_ssmh"]
    assert all("Invalid global border" in r["error"] for r in bad["particles"])
    ss = CompiledModel([tuple(d) for d in bad["program"]]).prog.subsampled
    assert ss[0]["kind"] == 2 and "Invalid global border" in ss[0]["reason"] and ss[1]["kind"] == 1
    # an observation that reads the principal node twice: the two local sections would share its nodes
    ss = CompiledModel([("assume", "mu", ["normal", 0, 1]), ("observe", ["normal", ["mul", "mu", "mu"], 1.0], 0.5),
                        ("observe", ["normal", ["mul", "mu", "mu"], 1.0], 0.7)]).prog.subsampled
    assert ss[0]["kind"] == 2 and "overlap" in ss[0]["reason"]


def test_node_counts_follow_the_scaffold_rules():
    # scaffold.py:201-218: the site, lookups of resampled nodes, deterministic applications downstream; a variadic
    # add is ONE application; an assume referenced by symbol is ONE node with one lookup per reference
    m = CompiledModel([("assume", "a", ["normal", 0, 1]), ("assume", "b", ["normal", 0, 1]), ("assume", "s", ["gamma", 1, 1])] +
                      [("observe", ["normal", ["add", ["mul", "a", float(i)], "b"], "s"], 2.0 * i) for i in range(5)])
    assert m.prog.affected == [16.0, 11.0, 6.0] and m.prog.n_groups == 1
    m = CompiledModel([("assume", "x", ["normal", 0, 1]), ("assume", "h", ["mul", "x", 2.0]),
                       ("observe", ["normal", "h", 1.0], 0.5), ("observe", ["normal", "h", 1.0], 0.7),
                       ("observe", ["normal", ["add", "h", "x", 1.0], 1.0], 0.1)])
    # x; lookups of x: in h's expression and in the add; h itself; lookups of h: 3; the add
    assert m.prog.affected == [1 + 2 + 1 + 3 + 1]
    assert m.prog.n_groups == 2 and [len(g.rows) for g in m.compiler.group_list] == [2, 1]


def test_blob_is_self_consistent():
    case = golden("program_chains.json")["hierarchy"]
    m = CompiledModel([tuple(d) for d in case["program"]])
    code, consts, meta, data, affected = m.prog.blob()
    assert code.dtype == np.int32 and code.size % compiler.INSTR_WORDS == 0 and meta.dtype == np.int32
    S, G, n_regs, n_sec = [int(v) for v in meta[:4]]
    assert S == 5 and len(affected) == 2 * S and n_regs <= compiler.MAX_REGS     # regular scaffolds, then the global sections
    n_roots = sum(len(ss["roots"]) for ss in m.prog.subsampled)
    n_cand = sum(len(cd) for cd in m.prog.candidates)
    tail = 3 * S + 1 + 2 * n_roots + 1 + 2 + S + 2 * S + n_cand     # subsampled-MH block, the scopes (only the default one), candidates
    directory = meta[-tail - 2 * n_sec: -tail].reshape(n_sec, 2)
    assert directory[:, 1].sum() == code.size // compiler.INSTR_WORDS
    ops = set(code.reshape(-1, compiler.INSTR_WORDS)[:, 0].tolist())
    assert ops <= set(compiler.OP.values())
    regs = code.reshape(-1, compiler.INSTR_WORDS)[:, 1]
    assert regs.min() >= 0 and regs.max() < n_regs
    assert sum(len(g.rows) for g in m.compiler.group_list) == m.n_observations == 6


def test_programs_outside_the_subset_are_refused():
    for prog, msg in [
        ([("assume", "x", ["normal", 0, 1]), ("assume", "f", ["lambda", ["y"], "y"])], "SP"),
        ([("assume", "x", ["normal", 0, 1]), ("observe", ["normal", ["normal", "x", 1.0], 1.0], 0.0)], "latent random choice"),
        ([("assume", "x", ["normal", "y", 1])], "Cannot find symbol"),
        ([("assume", "x", ["add", 1, 2])], "no unobserved random choice"),
        ([("assume", "x", ["normal", 0, 1]), ("observe", ["add", "x", 1.0], 0.0)], "random SPs can be observed"),
        ([("assume", "x", ["poisson", 3.0])], "outside"),
    ]:
        with pytest.raises(VentureException) as e:
            CompiledModel(prog)
        assert msg in str(e.value), (prog, str(e.value))


def test_compiled_model_speaks_the_interchange_format():
    """the directives of a compiled model are its program(): device particles restore into Lite traces (N3)"""
    m = CompiledModel([("assume", "p", ["beta", 1.0, 1.0]), ("observe", ["flip", "p"], True), ("observe", ["flip", "p"], False)])
    prog = m.program([])
    assert [d[0] for d in prog] == ["define", "observe", "observe"]
    vals = m.random_values([], np.array([0.25]))
    assert [v["value"] for v in vals] == [0.25, True, False]
    np.testing.assert_array_equal(m.sites_from_random_values([], vals), [0.25])
    assert oracle is not None


def test_check_applicability_follows_the_reference():
    """SubsampledBlockScaffoldIndexer.checkApplicability (infer/subsampled_mh.py:222-262): which warning, for which block.
    The fixture holds what Lite itself warned on the three programs (tests/golden/make_program_chains.py)."""
    case = golden("program_chains.json")
    lite = case["check_applicability"]
    good = CompiledModel([tuple(d) for d in case["normal_observes_ssmh"]["program"]])
    assert good.subsampled_applicability("one") == (True, None, "") and lite["normal_observes"] == []
    bad = CompiledModel([tuple(d) for d in case["invalid_border_ssmh"]["program"]])
    ok, kind, msg = bad.subsampled_applicability("one")
    assert not ok and [kind, "Invalid global border."] == lite["invalid_border"][0] and "Invalid global border" in msg
    # `scaled`: mu and s have borders, lo and x have none.  Lite stops at mu with "The scaffold section is not consistent
    # with the entire scaffold": its OrderedSet.__eq__ (backend/lite/orderedset.py:83-91) also compares the ORDER of the
    # two node sets, which here hold the same five nodes in different orders -- while subsampled_mh on that block runs
    # and is pinned decision for decision above.  The compiler reports what the structure says: no border at `lo`.
    m = CompiledModel([tuple(d) for d in case["scaled_ssmh"]["program"]])
    assert lite["scaled"] == [["SubsampledScaffoldNotApplicableWarning", "The scaffold section is not consistent with the entire scaffold."]]
    assert m.subsampled_applicability(0) == (True, None, "") and m.subsampled_applicability(1)[0]
    ok, kind, msg = m.subsampled_applicability("one")
    assert not ok and kind == "SubsampledScaffoldNotEffectiveWarning" and "same as mh" in msg
    assert m.subsampled_global_nodes("one") == 3.0 + 1.0 and m.subsampled_global_nodes(0) == 3.0


def blr_program(c, n_obs=None):
    """the logistic-regression program of SURVEY 8(d) D2 as directives (what tests/golden/make_lite_chains.py ran on Lite)"""
    D, N = c["D"], c["N"] if n_obs is None else n_obs
    return [("assume", "w_%d" % j, ["normal", 0.0, 1.0]) for j in range(D)] + \
           [("assume", "w", ["array"] + ["w_%d" % j for j in range(D)])] + \
           [("observe", ["log_odds_flip", ["vector_dot", "w", ["array"] + [float(v) for v in c["X"][i]]]], bool(c["y"][i]))
            for i in range(N)]


def test_compiled_logistic_regression_reproduces_lite():
    """`(array ...)` / `vector_dot` in the compiler: the C2 program text itself, compiled, against the Lite chains that
    pin the hand-written kernels (lite_chains.json c2): likelihood_weight, the increment, `one` and fixed-block MH"""
    c = golden("lite_chains.json")["c2"]
    D, N, P, K, seed = c["D"], c["N"], c["P"], c["K"], c["seed"]
    h = N // 2
    first, full = CompiledModel(blr_program(c, h)), CompiledModel(blr_program(c))
    assert full.prog.affected == c["untagged_nodes_one"][:1] * D and full.prog.n_groups == 1
    recs = c["particles"]
    vm = VM(first.prog.blob(), compiler.OP, P, seed)
    vm.step = c["steps"]["likelihood_weight"]
    vm.likelihood_weight()
    np.testing.assert_allclose(vm.lw, [unj(r["lw"]) for r in recs], rtol=RTOL)
    np.testing.assert_allclose(vm.sites.T, [r["w_lw"] for r in recs], rtol=RTOL)
    vm2 = VM(full.prog.blob(), compiler.OP, P, seed)
    vm2.sites[:] = vm.sites
    np.testing.assert_allclose(vm2.log_likelihood() - vm.log_likelihood(), [unj(r["incr"]) for r in recs], rtol=1e-10)
    vm2.step = c["steps"]["mh_one"]
    vm2.mh(K)
    np.testing.assert_allclose(vm2.sites.T, [r["w_one"] for r in recs], rtol=RTOL)
    np.testing.assert_allclose(vm2.log_likelihood(), [unj(r["ll_one"]) for r in recs], rtol=RTOL)
    for k in range(2 * D):                       # (resimulation_mh wts <k mod D> 1), one Philox step each
        vm2.step = c["steps"]["mh_each"] + k
        vm2.mh(1, block=k % D)
    np.testing.assert_allclose(vm2.sites.T, [r["w_each"] for r in recs], rtol=RTOL)


def test_compiled_logistic_regression_subsampled_mh_reproduces_lite():
    """the C5 program text compiled: the global border is the `w` array node, the local sections are the observations;
    every decision, section count, t statistic and the node count 3 of lite_chains.json c5"""
    c = golden("lite_chains.json")["c5"]
    m = CompiledModel(blr_program(c))
    assert [(s["kind"], len(s["roots"]), s["nodes"]) for s in m.prog.subsampled] == [(1, c["N"], 3.0)] * c["D"]
    recs = c["particles"]
    vm = VM(m.prog.blob(), compiler.OP, c["P"], c["seed"])
    vm.step = c["steps"]["prior"]
    vm.reset_to_prior()
    np.testing.assert_allclose(vm.sites.T, [r["w_prior"] for r in recs], rtol=RTOL)
    vm.step = c["steps"]["ssmh"]
    acc, visited, ts, nodes = [], [], [], []
    for _ in range(c["K"]):
        a, n, t, nd = vm.subsampled_mh(1, c["Nbatch"], c["k0"], c["eps"])
        acc.append(a), visited.append(n), ts.append(t), nodes.append(nd)
    acc, visited, ts, nodes = [np.array(v).T for v in (acc, visited, ts, nodes)]
    for p, r in enumerate(recs):
        np.testing.assert_array_equal(visited[p], r["n"])
        np.testing.assert_array_equal(acc[p] > 0, r["accept"])
        np.testing.assert_array_equal(nodes[p], r["nodes"])
        want = np.array([unj(v) for v in r["tstat"]])
        fin = np.isfinite(want)
        np.testing.assert_allclose(ts[p][fin], want[fin], rtol=1e-9)
    np.testing.assert_allclose(vm.sites.T, [r["w"] for r in recs], rtol=RTOL)


def tagged_blr_program(c, n_obs=None):
    """blr_model of tests/golden/make_lite_chains.py: every weight tagged into scope `wts`, block j"""
    prog = blr_program(c, n_obs)
    return [("assume", d[1], ["tag", ["quote", "wts"], float(j), d[2]]) if j < c["D"] else d for j, d in enumerate(prog)]


def test_tagged_scopes_reproduce_lite():
    """`(tag (quote wts) j (normal 0 1))`: one more deterministic node per weight in the scaffold (32 against 31), and
    `(resimulation_mh wts <block> 1)` picks the tagged site -- the model lite_chains.json c2 actually ran"""
    c = golden("lite_chains.json")["c2"]
    D, P, K, seed = c["D"], c["P"], c["K"], c["seed"]
    m = CompiledModel(tagged_blr_program(c))
    assert m.scopes == {"default": 0, "wts": 1} and m.scope_blocks == {1: dict((float(j), j) for j in range(D))}
    assert m.site_names == ["w_%d" % j for j in range(D)]
    recs = c["particles"]
    assert m.prog.affected == [recs[0]["nodes_one"][0]] * D == [32.0] * D
    vm = VM(m.prog.blob(), compiler.OP, P, seed)
    vm.sites[:] = np.array([r["w_lw"] for r in recs]).T
    vm.step = c["steps"]["mh_one"]
    _, nodes = vm.mh(K)
    np.testing.assert_allclose(vm.sites.T, [r["w_one"] for r in recs], rtol=RTOL)
    np.testing.assert_array_equal(nodes, [np.mean(r["nodes_one"]) for r in recs])
    for k in range(2 * D):
        vm.step = c["steps"]["mh_each"] + k
        _, nodes = vm.mh(1, block=m.scope_blocks[1][float(k % D)], scope=1)
        np.testing.assert_array_equal(nodes, [r["nodes_each"][k] for r in recs])
    np.testing.assert_allclose(vm.sites.T, [r["w_each"] for r in recs], rtol=RTOL)
    # a scope over a subset of the sites: `one` and `each` stay inside it
    sub = CompiledModel([("assume", "a", ["normal", 0.0, 1.0]), ("assume", "b", ["tag", ["quote", "s"], ["quote", "x"], ["normal", "a", 1.0]]),
                         ("assume", "c", ["tag", ["quote", "s"], 7, ["normal", "b", 1.0]]), ("observe", ["normal", "c", 1.0], 0.4)])
    assert sub.scope_blocks == {1: {"x": 0, 7.0: 1}} and sub.prog.scope_sites == [[0, 1, 2], [1, 2]]
    vm = VM(sub.prog.blob(), compiler.OP, 6, 3)
    vm.reset_to_prior()
    a0 = vm.sites[0].copy()
    vm.mh(40, block=-1, scope=1), vm.mh(4, block=-3, scope=1)
    np.testing.assert_array_equal(vm.sites[0], a0)
    with pytest.raises(VentureException) as e:
        CompiledModel([("assume", "a", ["tag", ["quote", "s"], 0, ["add", ["normal", 0.0, 1.0], ["normal", 0.0, 1.0]]])])
    assert "blocks of one random choice" in str(e.value)


@pytest.mark.parametrize("name", ["mixture_gibbs", "mixture_emap", "mixture3_gibbs"])
def test_compiled_gibbs_reproduces_lite(name):
    """(gibbs z one 1) x K and (gibbs z <block> 1) on Lite (EnumerativeGibbsOperator, infer/egibbs.py:42-107; emap:
    EnumerativeMAPOperator :117-121) vs the compiled program: a mixture with `biplex`, the three flip SPs and a flip whose
    probability hangs on another flip"""
    case = golden("program_chains.json")[name]
    is_map = name.endswith("emap")
    m = CompiledModel([tuple(d) for d in case["program"]])
    recs = case["particles"]
    assert m.site_names == recs[0]["sites"] and m.scopes == {"default": 0, "z": 1}
    vm = VM(m.prog.blob(), compiler.OP, case["P"], case["seed"])
    vm.step = case["steps"]["prior"]
    vm.reset_to_prior()
    np.testing.assert_allclose(vm.sites.T, [r["values_prior"] for r in recs], rtol=RTOL)
    vm.step = case["steps"]["gibbs_one"]
    nodes = []
    for _ in range(case["K"]):
        nodes.append(vm.gibbs(1, block=-1, scope=1, map=is_map)[1])
    np.testing.assert_array_equal(np.array(nodes).T, [r["nodes_one"] for r in recs])
    np.testing.assert_array_equal(vm.sites.T, [r["values_one"] for r in recs])
    nodes = []
    for k in range(2 * case["n_blocks"]):
        vm.step = case["steps"]["gibbs_block"] + k
        nodes.append(vm.gibbs(1, block=k % case["n_blocks"], scope=1, map=is_map)[1])
    np.testing.assert_array_equal(np.array(nodes).T, [r["nodes_block"] for r in recs])
    np.testing.assert_array_equal(vm.sites.T, [r["values_block"] for r in recs])
    np.testing.assert_allclose(vm.log_likelihood(joint=True), [unj(r["lj_at"]) for r in recs], rtol=RTOL)
