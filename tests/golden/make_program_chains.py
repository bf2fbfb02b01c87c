#!/usr/bin/env python
"""Generates tests/golden/program_chains.json: programs for the MODEL COMPILER (venturecxx_b200/compiler.py,
SURVEY 8(f) N2) run on THE REFERENCE'S LITE BACKEND in this container, with the machinery of
make_lite_chains.py -- Lite's own constructScaffold / detachAndExtract / regenAndAttach / mixMH on stand-in RNG
objects that return the draws the oracle's samplers make at the matching position of the Philox stream
(seed, particle, step).  The fixture holds the programs (as nested lists) and what Lite ended with: values,
weights, increments, node counts, log likelihoods; tests compile the same programs for the GPU and demand
equality.

  python oracle/lite_py3/build_lite_py3.py && python tests/golden/make_program_chains.py
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_lite_chains import Model, fl, install, mh_steps  # noqa: E402  (also puts oracle/_ref on sys.path)
import venture.lite.value as vv  # noqa: E402
import venture.value.dicts as vd  # noqa: E402


def programs():
    rng = np.random.default_rng(7)
    xs = np.round(rng.normal(size=9), 3).tolist()
    ys = np.round(1.5 * np.array(xs) - 0.7 + 0.4 * rng.normal(size=9), 3).tolist()
    regression = [("assume", "a", ["normal", 0.0, 2.0]), ("assume", "b", ["normal", 0.0, 2.0]),
                  ("assume", "s", ["gamma", 2.0, 2.0])] + \
                 [("observe", ["normal", ["add", ["mul", "a", x], "b"], "s"], y) for x, y in zip(xs, ys)]
    flips = [True, True, False, True, True, True, False, True, True, True, True, False]
    coin = [("assume", "p", ["beta", 1.5, 2.5])] + [("observe", ["flip", "p"], f) for f in flips]
    hier = [("assume", "mu", ["normal", 1.0, 3.0]), ("assume", "tau", ["gamma", 3.0, 2.0]),
            ("assume", "spread", ["div", 1.0, ["sqrt", "tau"]]),
            ("assume", "x1", ["normal", "mu", "spread"]), ("assume", "x2", ["normal", "mu", "spread"]),
            ("assume", "lo", ["uniform_continuous", -1.0, 0.5]),
            ("assume", "mid", ["add", "x1", "x2", "lo"]),
            ("observe", ["normal", "x1", 0.5], 2.1), ("observe", ["normal", "x2", 0.5], 0.3),
            ("observe", ["normal", ["mul", "mid", 0.5], 1.0], 1.0),
            ("observe", ["log_odds_flip", ["sub", "x1", "x2"]], True),
            ("observe", ["expon", ["exp", "lo"]], 0.8),
            ("observe", ["gamma", ["add", ["abs", "mu"], 1.0], "tau"], 1.7)]
    # real-valued sites only: Lite's slice evaluates the density outside a positive site's support and raises there
    regression2 = [("assume", "a", ["normal", 0.0, 2.0]), ("assume", "b", ["normal", 0.0, 2.0])] + \
                  [("observe", ["normal", ["add", ["mul", "a", x], "b"], 0.7], y) for x, y in zip(xs, ys)]
    chain = [("assume", "mu", ["normal", 1.0, 3.0]), ("assume", "x1", ["normal", "mu", 1.5]), ("assume", "x2", ["normal", "mu", 1.5]),
             ("observe", ["normal", "x1", 0.5], 2.1), ("observe", ["normal", "x2", 0.5], 0.3),
             ("observe", ["log_odds_flip", ["sub", "x1", "x2"]], True)]
    # subsampled MH: the reference's own pin model (test/inference_language/test_subsampledmh.py:87-99), smaller
    obs = np.round(rng.normal(10.0, 1.0, size=40), 3).tolist()
    normal_observes = [("assume", "mu", ["normal", 10.0, 3.0])] + [("observe", ["normal", "mu", 1.0], x) for x in obs]
    # border below the principal node (mu -> lookup -> mul = z), a latent child among the local sections, a second
    # site whose border is itself (s), and one without a border (lo: regular MH inside subsampledMixMH)
    cs = np.round(rng.normal(size=24), 2).tolist()
    zs = np.round(2.0 * 1.3 + np.array(cs) + 0.8 * rng.normal(size=24), 3).tolist()
    scaled = [("assume", "mu", ["normal", 1.0, 2.0]), ("assume", "s", ["gamma", 3.0, 3.0]),
              ("assume", "lo", ["uniform_continuous", -1.0, 1.0]),
              ("assume", "z", ["mul", 2.0, "mu"]), ("assume", "x", ["normal", "z", 1.5]),
              ("observe", ["normal", ["add", "lo", 0.3], 0.7], 0.2)] + \
             [("observe", ["normal", ["add", "z", c], "s"], v) for c, v in zip(cs, zs)]
    invalid = [("assume", "mu", ["normal", 0.0, 1.0]), ("assume", "x", ["normal", "mu", 1.0]),
               ("observe", ["normal", "x", 1.0], 0.5), ("observe", ["normal", "x", 1.0], 0.7)]
    # enumerative Gibbs on boolean sites: a two-component mixture with `biplex` (both branches evaluated, no brush), the
    # three flip SPs, a flip whose probability depends on another flip; the flips are tagged into scope `z`
    xs5 = [2.3, -0.8, 1.7, -1.4, 0.2]
    mixture = [("assume", "m0", ["normal", -1.0, 1.0]), ("assume", "m1", ["normal", 2.0, 1.0]),
               ("assume", "z0", ["tag", ["quote", "z"], 0, ["flip", 0.4]]),
               ("assume", "z1", ["tag", ["quote", "z"], 1, ["log_flip", -0.7]]),
               ("assume", "z2", ["tag", ["quote", "z"], 2, ["log_odds_flip", 0.3]]),
               ("assume", "z3", ["tag", ["quote", "z"], 3, ["flip", ["biplex", "z0", 0.8, 0.2]]]),
               ("assume", "z4", ["tag", ["quote", "z"], 4, ["flip", 0.5]])] + \
              [("observe", ["normal", ["biplex", "z%d" % i, "m1", "m0"], 0.7], x) for i, x in enumerate(xs5)]
    # categorical sites (weights literal and computed), `simplex`, `lookup` into an array of sites, literal objects
    xs3 = [0.1, 1.2, 2.2, -0.4]
    # (Lite's likelihood_weight regenerates the random choices in the order of its scaffold's border, which for this program is
    # not the order of the assumes: the test starts the chains from the recorded prior values instead of re-drawing them)
    mixture3 = [("assume", "m%d" % j, ["normal", float(j), 1.0]) for j in range(3)] + [("assume", "q", ["beta", 2.0, 2.0])] + \
               [("assume", "z%d" % i, ["categorical", ["simplex", 0.2, 0.5, 0.3]]) for i in range(3)] + \
               [("assume", "z3", ["categorical", ["simplex", "q", ["sub", 1.0, "q"]], ["array", 2.0, 0.0]])] + \
               [("observe", ["normal", ["lookup", ["array", "m0", "m1", "m2"], "z%d" % i], 0.7], x) for i, x in enumerate(xs3)]
    # the same with the assignments tagged into scope `z`, for enumerative gibbs over categorical sites (with explicit objects:
    # Lite's CategoricalOutputPSP.enumerateValues returns bare python ints for the default objects and fails to wrap them)
    mixture3z = [("assume", "m%d" % j, ["normal", float(j), 1.0]) for j in range(3)] + [("assume", "q", ["beta", 2.0, 2.0])] + \
                [("assume", "z%d" % i, ["tag", ["quote", "z"], i, ["categorical", ["simplex", 0.2, 0.5, 0.3], ["array", 0.0, 1.0, 2.0]]]) for i in range(3)] + \
                [("assume", "z3", ["tag", ["quote", "z"], 3, ["categorical", ["simplex", "q", ["sub", 1.0, "q"]], ["array", 2.0, 0.0]]])] + \
                [("observe", ["normal", ["lookup", ["array", "m0", "m1", "m2"], "z%d" % i], 0.7], x) for i, x in enumerate(xs3)]
    return {"mixture3z": mixture3z, "mixture3": mixture3, "mixture": mixture, "regression": regression, "coin": coin, "hierarchy": hier, "regression2": regression2, "chain": chain,
            "normal_observes": normal_observes, "scaled": scaled, "invalid_border": invalid}


def box(v):
    return vd.boolean(v) if isinstance(v, bool) else vd.number(float(v))


def run(name, prog, seed, P, K, n_first):
    """n_first observations are incorporated before likelihood_weight, the rest after (increments)."""
    out = {"seed": seed, "P": P, "K": K, "program": [list(d) for d in prog], "n_first": n_first,
           "steps": {"likelihood_weight": 4, "mh": 30}, "particles": []}
    assumes = [d for d in prog if d[0] == "assume"]
    observes = [d for d in prog if d[0] == "observe"]
    for p in range(P):
        m = Model()
        dids = [m.assume(d[1], d[2]) for d in assumes]
        for d in observes[:n_first]:
            m.observe(d[1], box(d[2]))
        m.t.makeConsistent()
        install(m, seed, p)
        # the default scope's blocks are the unconstrained random choices; canonical order = order of the assumes
        random_dids = [did for did, d in zip(dids, assumes) if m.t.families[did] in m.t.rcs]
        nodes = [m.t.families[did] for did in random_dids]
        m.tape.site_of = lambda key: [i for i, n in enumerate(nodes) if n is key][0]
        rec = {"sites": [d[1] for did, d in zip(dids, assumes) if did in random_dids]}
        m.tape.begin(4, 0)
        rec["lw"] = fl(m.t.likelihood_weight())
        rec["values_lw"] = [fl(m.value(did)) for did in random_dids]
        for d in observes[n_first:]:
            m.observe(d[1], box(d[2]))
        rec["incr"] = fl(m.t.makeConsistent())
        rec["nodes"] = mh_steps(m, 30, K)
        rec["values_mh"] = [fl(m.value(did)) for did in random_dids]
        rec["ll_at"] = fl(m.t.log_likelihood_at(vv.VentureSymbol("default"), vv.VentureSymbol("all")))
        rec["lj_at"] = fl(m.t.log_joint_at(vv.VentureSymbol("default"), vv.VentureSymbol("all")))
        out["particles"].append(rec)
    return out


def run_slice(prog, seed, P, K, w, m, op="slice"):
    """(slice default one w m 1) K times: draw 0 = block choice, then slice_sample.py's uniforms in call order"""
    from make_lite_chains import infer_exp
    out = {"seed": seed, "P": P, "K": K, "w": w, "m": m, "program": [list(d) for d in prog], "steps": {"slice": 50}, "particles": []}
    assumes = [d for d in prog if d[0] == "assume"]
    observes = [d for d in prog if d[0] == "observe"]
    for p in range(P):
        mdl = Model()
        dids = [mdl.assume(d[1], d[2]) for d in assumes]
        for d in observes:
            mdl.observe(d[1], box(d[2]))
        mdl.t.makeConsistent()
        install(mdl, seed, p)
        random_dids = [did for did, d in zip(dids, assumes) if mdl.t.families[did] in mdl.t.rcs]
        nodes = [mdl.t.families[did] for did in random_dids]
        mdl.tape.site_of = lambda key: [i for i, n in enumerate(nodes) if n is key][0]
        rec = {"values_start": [fl(mdl.value(did)) for did in random_dids], "nodes": []}
        mdl.tape.sequential = True
        for k in range(K):
            mdl.tape.begin(50 + k, 1)
            rec["nodes"].append(mdl.t.primitive_infer(infer_exp(op, "default", "one", w, m, 1)))
        mdl.tape.sequential = False
        rec["values_slice"] = [fl(mdl.value(did)) for did in random_dids]
        rec["lj_at"] = fl(mdl.t.log_joint_at(vv.VentureSymbol("default"), vv.VentureSymbol("all")))
        out["particles"].append(rec)
    return out


def run_subsampled(prog, seed, P, K, Nbatch, k0, eps, block="one", make_consistent=False):
    """(subsampled_mh default <block> Nbatch k0 eps false 0 false 1) K times on Lite (infer/subsampled_mh.py), np_rng.permutation
    standing in as the kernels' Feistel permutation of (seed, step); records sequentialTest's own return values.
    make_consistent: SubsampledMHOperator().makeConsistent over every block (what `subsampled_mh_make_consistent default one`
    is meant to run; its dispatch entry, infer/dispatch.py:158-161, reads an unbound `extra`) after every transition -- an accepted
    proposal leaves the unvisited local sections stale in Lite (:57-72), and ANOTHER site's transition would then read
    stale parents; the device keeps no per-section cache, i.e. it is Lite with the sections made consistent."""
    import ctypes
    import oracle
    import venture.lite.infer.subsampled_mh as ss
    out = {"seed": seed, "P": P, "K": K, "Nbatch": Nbatch, "k0": k0, "eps": eps, "block": block,
           "program": [list(d) for d in prog], "steps": {"prior": 0, "ssmh": 9}, "particles": []}
    assumes = [d for d in prog if d[0] == "assume"]
    observes = [d for d in prog if d[0] == "observe"]
    tests = []
    orig = ss.sequentialTest

    def recording(*a):
        r = orig(*a)
        tests.append(r)
        return r
    ss.sequentialTest = recording
    for p in range(P):
        mdl = Model()
        install(mdl, seed, p)
        mdl.tape.begin(0, 0)
        dids = [mdl.assume(d[1], d[2]) for d in assumes]
        for d in observes:
            mdl.observe(d[1], box(d[2]))
        mdl.t.makeConsistent()
        random_dids = [did for did, d in zip(dids, assumes) if mdl.t.families[did] in mdl.t.rcs]
        nodes = [mdl.t.families[did] for did in random_dids]
        tape = mdl.tape
        tape.site_of = lambda key: [i for i, n in enumerate(nodes) if n is key][0]

        def permutation(items, tape=tape):
            n = len(items)
            perm = np.empty(n, dtype=np.uint64)
            oracle.lib().orc_feistel_perm(ctypes.c_uint64(tape.seed), ctypes.c_uint32(tape.step), ctypes.c_uint64(n),
                                          perm.ctypes.data_as(ctypes.c_void_p))
            return [items[int(q)] for q in perm]
        mdl.t.np_rng.permutation = permutation
        rec = {"values_prior": [fl(mdl.value(did)) for did in random_dids]}
        del tests[:]
        from make_lite_chains import infer_exp
        rec["nodes"] = []
        try:
            for k in range(K):
                mdl.tape.begin(9 + k, 2)
                rec["nodes"].append(mdl.t.primitive_infer(infer_exp("subsampled_mh", "default", block, Nbatch, k0, eps, False, 0, False, 1)))
                if make_consistent:
                    sc, blk = mdl.t._normalizeEvaluatedScopeAndBlock(vv.VentureSymbol("default"), vv.VentureSymbol("one"))
                    ss.SubsampledMHOperator().makeConsistent(mdl.t, ss.SubsampledBlockScaffoldIndexer(
                        sc, blk, useDeltaKernels=False, deltaKernelArgs=0, updateValues=False))
        except ss.SubsampledScaffoldError as e:
            rec["error"] = str(e)
        rec["accept"] = [bool(t[0]) for t in tests]
        rec["n"] = [float(t[1]) for t in tests]
        rec["tstat"] = [fl(t[2]) for t in tests]
        rec["values"] = [fl(mdl.value(did)) for did in random_dids]
        out["particles"].append(rec)
    ss.sequentialTest = orig
    return out


def site_names_of(prog):
    """the assumes that are random choices (possibly under a tag), by the compiler's own account"""
    sys.path.insert(0, os.path.join(HERE, "..", ".."))
    from venturecxx_b200.compiler import CompiledModel
    return CompiledModel(prog).site_names


def run_gibbs(prog, seed, P, K, scope, n_blocks, op="gibbs"):
    """(gibbs <scope> one 1) K times, then (gibbs <scope> <block> 1) for every block in order: EnumerativeGibbsOperator,
    infer/egibbs.py:42-107; draw 0 block choice, draw 1 the accept uniform, draw 2 the categorical over the candidates"""
    from make_lite_chains import infer_exp
    out = {"seed": seed, "P": P, "K": K, "scope": scope, "n_blocks": n_blocks, "program": [list(d) for d in prog],
           "steps": {"prior": 0, "gibbs_one": 20, "gibbs_block": 200}, "particles": []}
    assumes = [d for d in prog if d[0] == "assume"]
    observes = [d for d in prog if d[0] == "observe"]
    for p in range(P):
        mdl = Model()
        install(mdl, seed, p)
        mdl.tape.begin(0, 0)
        mdl.tape.sequential = True             # the flips' priors draw from py_rng.random(): in call order here
        dids = [mdl.assume(d[1], d[2]) for d in assumes]
        mdl.tape.sequential = False            # from now on py_rng.random() is mixMH's accept uniform (draw 1)
        for d in observes:
            mdl.observe(d[1], box(d[2]))
        mdl.t.makeConsistent()
        site_names = set(site_names_of(prog))
        random_dids = [did for did, d in zip(dids, assumes) if d[1] in site_names]
        mdl.tape.site_of = lambda key: int(key.getNumber())          # blocks 0..n-1 of the scope, in order

        def values():
            return [fl(float(mdl.value(did))) if not isinstance(mdl.value(did), bool) else (1.0 if mdl.value(did) else 0.0) for did in random_dids]
        rec = {"sites": [d[1] for did, d in zip(dids, assumes) if did in random_dids], "values_prior": values(), "nodes_one": [], "nodes_block": []}
        for k in range(K):
            mdl.tape.begin(20 + k, 2)
            rec["nodes_one"].append(mdl.t.primitive_infer(infer_exp(op, scope, "one", 1)))
        rec["values_one"] = values()
        for k in range(2 * n_blocks):
            mdl.tape.begin(200 + k, 2)
            rec["nodes_block"].append(mdl.t.primitive_infer(infer_exp(op, scope, k % n_blocks, 1)))
        rec["values_block"] = values()
        rec["lj_at"] = fl(mdl.t.log_joint_at(vv.VentureSymbol("default"), vv.VentureSymbol("all")))
        out["particles"].append(rec)
    return out


def run_check_applicability(prog):
    """(subsampled_mh_check_applicability default one) on Lite: the warning classes it raises, in order"""
    import warnings
    from make_lite_chains import infer_exp
    assumes = [d for d in prog if d[0] == "assume"]
    observes = [d for d in prog if d[0] == "observe"]
    mdl = Model()
    install(mdl, 3, 0)
    mdl.tape.begin(0, 0)
    for d in assumes:
        mdl.assume(d[1], d[2])
    for d in observes:
        mdl.observe(d[1], box(d[2]))
    mdl.t.makeConsistent()
    with warnings.catch_warnings(record=True) as caught:
        warnings.simplefilter("always")
        mdl.t.primitive_infer(infer_exp("subsampled_mh_check_applicability", "default", "one", 1))
    return [[type(w.message).__name__, str(w.message)] for w in caught if "Subsampled" in type(w.message).__name__]


def main():
    fixture = {}
    progs = programs()
    fixture["regression"] = run("regression", progs["regression"], 23, 5, 40, 4)
    fixture["coin"] = run("coin", progs["coin"], 29, 5, 30, 6)
    fixture["hierarchy"] = run("hierarchy", progs["hierarchy"], 31, 5, 60, 3)
    fixture["regression_slice"] = run_slice(progs["regression2"], 37, 4, 30, 0.7, 8)
    fixture["chain_slice"] = run_slice(progs["chain"], 41, 4, 40, 1.3, 5)
    # (slice_doubling default one w p 1): DoublingSliceOperator, infer/slice_sample.py:148-196
    fixture["regression_slice_doubling"] = run_slice(progs["regression2"], 59, 4, 30, 0.4, 5, op="slice_doubling")
    fixture["chain_slice_doubling"] = run_slice(progs["chain"], 61, 4, 40, 0.9, 3, op="slice_doubling")
    fixture["mixture3"] = run("mixture3", progs["mixture3"], 73, 5, 80, 4)
    fixture["mixture_gibbs"] = run_gibbs(progs["mixture"], 67, 5, 40, "z", 5)
    fixture["mixture3_gibbs"] = run_gibbs(progs["mixture3z"], 79, 5, 40, "z", 4)
    fixture["mixture_emap"] = run_gibbs(progs["mixture"], 71, 5, 12, "z", 5, op="emap")      # EnumerativeMAPOperator, egibbs.py:117-121
    fixture["normal_observes_ssmh"] = run_subsampled(progs["normal_observes"], 43, 5, 40, 5, 2, 0.05)
    fixture["scaled_ssmh"] = run_subsampled(progs["scaled"], 47, 5, 60, 4, 2, 0.1, make_consistent=True)
    fixture["invalid_border_ssmh"] = run_subsampled(progs["invalid_border"], 53, 2, 8, 1, 1, 0.1)
    fixture["check_applicability"] = dict((k, run_check_applicability(progs[k])) for k in ("normal_observes", "scaled", "invalid_border"))
    fixture["_source"] = ("reference Lite backend (py3 rendering by oracle/lite_py3/build_lite_py3.py) run by "
                          "tests/golden/make_program_chains.py; numpy %s" % np.__version__)
    path = os.path.join(HERE, "program_chains.json")
    with open(path, "w") as f:
        json.dump(fixture, f, indent=1, sort_keys=True)
    print("wrote", path)


if __name__ == "__main__":
    main()
