#!/usr/bin/env python
"""Generates tests/golden/lite_chains.json by RUNNING THE REFERENCE'S LITE BACKEND (its own
Trace / constructScaffold / detachAndExtract / regenAndAttach / mixMH / MHOperator /
EnumerativeGibbsOperator / subsampledMixMH code) in this container.

  python oracle/lite_py3/build_lite_py3.py      # py3 rendering of backend/lite -> oracle/_ref/ (git-ignored)
  python tests/golden/make_lite_chains.py       # runs Lite, writes the fixture

How the chains are made comparable.  Lite draws from numpy's MT19937 and Python's `random`; the
kernels (and oracle/oracle.c) draw from Philox streams addressed by (seed, particle, step, draw).
Sampler streams cannot agree, so each Lite trace gets stand-in `np_rng` / `py_rng` objects whose
methods return the draw the ORACLE's sampler makes at the matching position of the matching Philox
stream (oracle.draw -> orc_draw).  Everything else -- which nodes are in the DRG, which absorb, the
order of detach and regen, every logDensity, (un)incorporate on the count tables, rho, xi, the
+-inf policy, the accept test, constrain / makeConsistent, likelihood_weight -- is the reference's
code, unmodified apart from the syntactic py3 rendering.  The committed fixture holds only inputs
and the states / weights Lite ended with; tests/test_oracle_lite_chains.py replays the same
(seed, particle, step) in oracle.c and demands equality (1e-12 relative for reals, exact for
discrete state), and tests/test_gpu_lite_chains.py does the same through the C ABI on the GPU.

Draw layout of one MH transition at `step` (the oracle's convention, see oracle.c):
  draw 0      block choice u_sel: block index = min(int(u_sel * n_blocks), n_blocks-1) in the model's
              canonical site order (Lite's SamplableMap order is relabelled to that order here --
              a uniform choice does not depend on labels)
  draw 1      accept uniform (mh.py:85)
  draw 2...   the proposal's value draws, in regen order
likelihood_weight / prior resimulation at `step`: value draws from draw 0 in regen order.
"""
import json
import math
import os
import sys
import warnings

warnings.simplefilter("ignore")
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))

import numpy as np  # noqa: E402

import oracle  # noqa: E402
import venture.lite.value as vv  # noqa: E402
import venture.value.dicts as vd  # noqa: E402
from venture.lite.trace import Trace  # noqa: E402


# ------------------------------------------------------------------ expressions
def sx(e):
    """Nested python lists / str / numbers -> the stack-dict expression Trace.eval takes
    (lite/trace.py:397-401), with the two macros the sivm would have expanded
    (python/lib/sivm/macro.py:293-311): lambda -> make_csp, if -> biplex."""
    if isinstance(e, str):
        return vd.symbol(e)
    if isinstance(e, bool):
        return vd.boolean(e)
    if isinstance(e, int):
        return vd.integer(e) if False else vd.number(e)
    if isinstance(e, float):
        return vd.number(e)
    if isinstance(e, dict):
        return e
    assert isinstance(e, list)
    if e and e[0] == "lambda":
        return [vd.symbol("make_csp"), vd.quote([sx(p) for p in e[1]]), vd.quote(sx(e[2]))]
    if e and e[0] == "if":
        return [[vd.symbol("biplex"), sx(e[1]), sx(["lambda", [], e[2]]), sx(["lambda", [], e[3]])]]
    if e and e[0] == "quote":
        return vd.quote(sx(e[1]))
    return [sx(x) for x in e]


class Model(object):
    """One Lite trace with directive bookkeeping (what engine/trace.py:53-126 does around it)."""

    def __init__(self, seed=1):
        self.t = Trace(seed)
        self.did = 0

    def assume(self, name, exp):
        self.did += 1
        self.t.eval(self.did, sx(exp))
        self.t.bindInGlobalEnv(name, self.did)
        return self.did

    def observe(self, exp, value):
        self.did += 1
        self.t.eval(self.did, sx(exp))
        self.t.observe(self.did, value)
        return self.did

    def value(self, did):
        return self.t.extractValue(did)["value"]


# ------------------------------------------------------------------ RNG stand-ins
class Tape(object):
    """Cursor into Philox stream (seed, particle, step) shared by the two stand-ins."""

    def __init__(self, seed, particle):
        self.seed, self.particle = seed, particle
        self.step, self.d = 0, 0
        self.site_of = None     # block key -> canonical site index
        self.log = []

    def begin(self, step, d0):
        self.step, self.d = step, d0

    def at(self, d):
        return oracle.uniforms(self.seed, self.particle, self.step, d + 1)[d]

    def draw(self, kind, params):
        v, self.d = oracle.draw(kind, params, self.seed, self.particle, self.step, self.d)
        return v


class PyRng(object):
    """Stands in for trace.py_rng (random.Random): block choice and the accept uniform."""

    def __init__(self, tape):
        self.tape = tape

    def sample(self, population, k):          # smap.py:56 <- trace.py:320 sampleBlock
        assert k == 1
        nb = len(population)
        b = min(int(self.tape.at(0) * nb), nb - 1)
        for entry in population:
            if self.tape.site_of(entry[0]) == b:
                return [entry]
        raise KeyError(b)

    def random(self):                          # mh.py:85
        if getattr(self.tape, "sequential", False):     # slice_sample.py draws its uniforms in call order
            return self.tape.draw(0, [0.0])
        return self.tape.at(1)

    def uniform(self, a, b):                   # random.uniform: a + (b - a) * random()
        return a + (b - a) * self.random()

    # FixedRandomness (utils.py:369-394) saves / restores the generator around evaluations of the slice density
    def getstate(self):
        return self.tape.d

    def setstate(self, d):
        self.tape.d = d

    def jumpahead(self, _n):
        pass

    def randint(self, lo, hi):                 # particle.py:97-99 seeds particle-local RNGs (egibbs); those
        return 12345                           # are never drawn from on these paths (a draw would raise here)


class NpRng(object):
    """Stands in for trace.np_rng (numpy RandomState): the value draws of regen."""

    def __init__(self, tape):
        self.tape = tape

    def normal(self, loc=0.0, scale=1.0):      # continuous.py:317
        return self.tape.draw(1, [loc, scale])

    def gamma(self, shape, scale=1.0):         # continuous.py:1164 (scale = 1/rate)
        return self.tape.draw(2, [shape, 1.0 / scale])

    def multinomial(self, n, ps):              # utils.py:104 simulateCategorical
        assert n == 1
        k = int(self.tape.draw(3, list(ps)))
        out = np.zeros(len(ps), dtype=int)
        out[k] = 1
        return out

    def random_sample(self):
        return self.tape.draw(0, [0.0])

    def get_state(self):                       # FixedRandomness, utils.py:381-393
        return self.tape.d

    def set_state(self, d):
        self.tape.d = d

    def seed(self, _s):
        pass

    def exponential(self, scale=1.0):            # continuous.py:1134 (inverse CDF, as the kernels' expon_sim)
        return -math.log1p(-self.tape.draw(0, [0.0])) * scale

    def uniform(self, low=0.0, high=1.0):        # continuous.py:637
        return low + (high - low) * self.tape.draw(0, [0.0])


def install(model, seed, particle):
    tape = Tape(seed, particle)
    model.t.py_rng = PyRng(tape)
    model.t.np_rng = NpRng(tape)
    model.tape = tape
    return tape


def infer_exp(op, scope, block, *extra):
    def lift(x):
        if isinstance(x, bool):
            return vv.VentureBool(x)
        if isinstance(x, (int, float)):
            return vv.VentureNumber(x)
        return vv.VentureSymbol(x)
    return [op, lift(scope), lift(block)] + [lift(x) for x in extra]


def mh_steps(model, step0, n_steps, block="one", op="resimulation_mh", extra=()):
    """n_steps transitions, one primitive_infer call each (transitions = 1) so the stand-ins can be
    pointed at Philox step step0 + k; returns Lite's own return values (mean node counts)."""
    out = []
    for k in range(n_steps):
        model.tape.begin(step0 + k, 2)
        out.append(model.t.primitive_infer(infer_exp(op, "default", block, *(tuple(extra) + (1,)))))
    return out


def fl(x):
    x = float(x)
    if math.isnan(x):
        return "nan"
    if math.isinf(x):
        return "inf" if x > 0 else "-inf"
    return x


# ------------------------------------------------------------------ C1 normal-normal
def chain_c1(seed=11, P=6, K=12):
    mu0, s0, s = 0.5, 2.0, 1.5
    ys = [2.0, 1.0, -0.5]
    out = {"seed": seed, "P": P, "K": K, "mu0": mu0, "s0": s0, "s": s, "y": ys,
           "steps": {"likelihood_weight": 3, "mh": 10}, "particles": []}
    for p in range(P):
        m = Model()
        x = m.assume("x", ["normal", mu0, s0])
        for y in ys[:2]:
            m.observe(["normal", "x", s], vd.number(y))
        m.t.makeConsistent()
        install(m, seed, p)
        m.tape.site_of = lambda key: 0
        rec = {}
        m.tape.begin(3, 0)
        rec["lw"] = fl(m.t.likelihood_weight())                       # lite/trace.py:522-530
        rec["x_lw"] = fl(m.value(x))
        m.observe(["normal", "x", s], vd.number(ys[2]))
        rec["incr"] = fl(m.t.makeConsistent())                        # lite/trace.py:422-448
        rec["nodes"] = mh_steps(m, 10, K)
        rec["x_mh"] = fl(m.value(x))
        rec["ll_at"] = fl(m.t.log_likelihood_at(vv.VentureSymbol("default"), vv.VentureSymbol("all")))
        rec["lj_at"] = fl(m.t.log_joint_at(vv.VentureSymbol("default"), vv.VentureSymbol("all")))
        out["particles"].append(rec)
    return out


# ------------------------------------------------------------------ C2 logistic regression
def blr_data(N, D, seed):
    X = np.random.default_rng(seed).normal(size=(N, D))
    w_true = np.random.default_rng(seed + 1).normal(size=D)
    pr = 1 / (1 + np.exp(-X @ w_true))
    y = (np.random.default_rng(seed + 2).random(N) < pr).astype(np.uint8)
    return X, y


def blr_model(D, X, y, n_first):
    m = Model()
    # the weights are also tagged into scope `wts`, block j, so that `each` and fixed-block programs
    # exist (the default scope admits only one/all/none/ordered, infer/mh.py:119-123)
    wd = [m.assume("w_%d" % j, ["tag", ["quote", "wts"], j, ["normal", 0.0, 1.0]]) for j in range(D)]
    m.assume("w", ["array"] + ["w_%d" % j for j in range(D)])
    for i in range(n_first):
        m.observe(["log_odds_flip", ["vector_dot", "w", ["array"] + [float(v) for v in X[i]]]], vd.boolean(bool(y[i])))
    return m, wd


def chain_c2(seed=17, P=4, D=5, N=14, K=24):
    X, y = blr_data(N, D, 40)
    h = N // 2
    out = {"seed": seed, "P": P, "D": D, "N": N, "K": K, "data_seed": 40, "X": X.tolist(), "y": y.tolist(),
           "steps": {"likelihood_weight": 5, "mh_one": 20, "mh_each": 100, "mh_all": 200}, "particles": []}
    for p in range(P):
        m, wd = blr_model(D, X, y, h)
        m.t.makeConsistent()
        install(m, seed, p)
        nodes = [list(m.t.scopes[vv.VentureSymbol("wts")][vv.VentureNumber(j)])[0] for j in range(D)]
        m.tape.site_of = lambda key: [i for i, n in enumerate(nodes) if n is key][0]
        rec = {}
        m.tape.begin(5, 0)
        rec["lw"] = fl(m.t.likelihood_weight())
        rec["w_lw"] = [fl(m.value(d)) for d in wd]
        for i in range(h, N):
            m.observe(["log_odds_flip", ["vector_dot", "w", ["array"] + [float(v) for v in X[i]]]], vd.boolean(bool(y[i])))
        rec["incr"] = fl(m.t.makeConsistent())
        rec["nodes_one"] = mh_steps(m, 20, K)
        rec["w_one"] = [fl(m.value(d)) for d in wd]
        rec["ll_one"] = fl(m.t.log_likelihood_at(vv.VentureSymbol("default"), vv.VentureSymbol("all")))
        # (resimulation_mh wts each 2) = every block in order, twice (dispatch.py:86-128), issued block by
        # block so that each transition sits on its own Philox step
        rec["nodes_each"] = []
        for k in range(2 * D):
            m.tape.begin(100 + k, 2)
            rec["nodes_each"].append(m.t.primitive_infer(infer_exp("resimulation_mh", "wts", k % D, 1)))
        rec["w_each"] = [fl(m.value(d)) for d in wd]
        rec["nodes_all"] = mh_steps(m, 200, 4, block="all")
        rec["w_all"] = [fl(m.value(d)) for d in wd]
        rec["lj_all"] = fl(m.t.log_joint_at(vv.VentureSymbol("default"), vv.VentureSymbol("all")))
        out["particles"].append(rec)
    # transloop's return value for the model exactly as SURVEY 8(d) D2 writes it (no tag wrapper)
    m = Model()
    wd = [m.assume("w_%d" % j, ["normal", 0.0, 1.0]) for j in range(D)]
    m.assume("w", ["array"] + ["w_%d" % j for j in range(D)])
    for i in range(N):
        m.observe(["log_odds_flip", ["vector_dot", "w", ["array"] + [float(v) for v in X[i]]]], vd.boolean(bool(y[i])))
    m.t.makeConsistent()
    install(m, seed, 0)
    nodes = [m.t.families[d] for d in wd]
    m.tape.site_of = lambda key: [i for i, n in enumerate(nodes) if n is key][0]
    out["untagged_nodes_one"] = mh_steps(m, 900, 2)
    out["untagged_nodes_all"] = mh_steps(m, 910, 2, block="all")
    return out


# ------------------------------------------------------------------ C3 HMM (examples/hmm.vnt:31-66)
def hmm_model():
    t = .33333333333333333
    m = Model()
    rows = [[t, t, t, 0, 0], [0, t, t, t, 0], [0, 0, t, t, t], [t, 0, 0, t, t], [t, t, 0, 0, t], [t, t, t, 0, 0]]
    erows = [[.6 if i == j else .1 for j in range(5)] for i in range(5)]
    m.assume("transition_matrix", ["array"] + [["simplex"] + [float(x) for x in r] for r in rows])
    m.assume("emission_matrix", ["array"] + [["simplex"] + r for r in erows])
    objs = ["array", 1, 2, 3, 4, 5]
    m.assume("transition", ["lambda", ["state"], ["categorical", ["lookup", "transition_matrix", ["sub", "state", 1]], objs]])
    m.assume("emission", ["lambda", ["state"], ["categorical", ["lookup", "emission_matrix", ["sub", "state", 1]], objs]])
    m.assume("state", ["mem", ["lambda", ["t"], ["tag", ["quote", "seq"], "t",
                                                  ["if", ["lte", "t", 1], 1, ["transition", ["state", ["sub", "t", 1]]]]]]])
    return m, rows, erows


def chain_c3(seed=23, P=6, K=40):
    data = [3, 2, 3, 5, 5, 1, 2, 5, 2, 4, 4, 1, 2, 2, 1, 4, 2, 3, 5, 1]
    T = len(data)
    out = {"seed": seed, "P": P, "K": K, "T": T, "data": data, "steps": {"filter": 1, "mh_one": 50, "mh_each": 200},
           "particles": []}
    for p in range(P):
        m, rows, erows = hmm_model()
        install(m, seed, p)
        rec = {"incr": []}
        sd = []
        for t in range(1, T + 1):
            m.tape.begin(1 + (t - 1), 0)     # evaluating state(t) simulates the transition at draw 0 (regen.py:255-285)
            m.observe(["emission", ["state", t]], vd.number(data[t - 1]))
            rec["incr"].append(fl(m.t.makeConsistent()))
        nodes = []
        for t in range(2, T + 1):
            pn = list(m.t.getNodesInBlock(vv.VentureSymbol("seq"), vv.VentureNumber(t)))
            assert len(pn) == 1
            nodes.append(pn[0])

        def hist():
            return [int(m.t.valueAt(n).getNumber()) for n in nodes]
        rec["hist_filter"] = [1] + hist()
        m.tape.site_of = lambda key: [i for i, n in enumerate(nodes) if n is key][0]
        rec["nodes_one"] = mh_steps(m, 50, K)
        rec["hist_one"] = [1] + hist()
        rec["nodes_each"] = []
        for k in range(2 * (T - 1)):
            m.tape.begin(200 + k, 2)
            rec["nodes_each"].append(m.t.primitive_infer(infer_exp("resimulation_mh", "seq", 2 + k % (T - 1), 1)))
        rec["hist_each"] = [1] + hist()
        rec["ll_at"] = fl(m.t.log_likelihood_at(vv.VentureSymbol("default"), vv.VentureSymbol("all")))
        rec["lj_at"] = fl(m.t.log_joint_at(vv.VentureSymbol("default"), vv.VentureSymbol("all")))
        # enumerative Gibbs (infer/egibbs.py:42-107): (gibbs default one 1) x K, then two ordered sweeps
        rec["nodes_gibbs_one"] = mh_steps(m, 300, K, op="gibbs")
        rec["hist_gibbs_one"] = [1] + hist()
        rec["nodes_gibbs_each"] = []
        for k in range(2 * (T - 1)):
            m.tape.begin(400 + k, 2)
            rec["nodes_gibbs_each"].append(m.t.primitive_infer(infer_exp("gibbs", "seq", 2 + k % (T - 1), 1)))
        rec["hist_gibbs_each"] = [1] + hist()
        out["particles"].append(rec)
    out["steps"]["gibbs_one"], out["steps"]["gibbs_each"] = 300, 400
    out["trans"], out["emit"] = rows, erows
    return out


# ------------------------------------------------------------------ C4 LDA (examples/lda.vnt:18-41)
def lda_model(K, V, seed, particle, prior_step, tag_z=False):
    m = Model()
    install(m, seed, particle)
    m.tape.begin(prior_step, 0)       # the two gamma assumes evaluate on draws 0.. of `prior_step`, in order
    m.assume("n_topics", K)
    m.assume("vocab_size", V)
    ad = m.assume("alpha_document_topic", ["gamma", 1.0, 2.0])
    aw = m.assume("alpha_topic_word", ["gamma", 1.0, 2.0])
    m.assume("doc", ["mem", ["lambda", ["doc_id"], ["make_sym_dir_cat", "alpha_document_topic", "n_topics"]]])
    m.assume("topic", ["mem", ["lambda", ["topic_id"], ["make_sym_dir_cat", "alpha_topic_word", "vocab_size"]]])
    zexp = [["doc", "doc_id"]]
    if tag_z:   # lda.vnt with the topic draw tagged into scope `z`, block 1000 doc_id + pos
        zexp = ["tag", ["quote", "z"], ["add", ["mul", "doc_id", 1000], "pos"], zexp]
    m.assume("word", ["mem", ["lambda", ["doc_id", "pos"], [["topic", zexp]]]])
    return m, ad, aw


def patch_dircat_sampler(K, V):
    """CDirCatOutputPSP.simulate (dirichlet.py:164-166) descends a sum tree with one uniform per level;
    the kernels use one uniform and the inverse CDF over (alpha + counts).  Same law, different map from
    uniforms to outcomes, so the sampler is replaced by the oracle's (orc_draw kind 4) like every other
    sampler here.  The token application ((topic z)) is re-simulated on every regen and immediately
    re-constrained (regen.py:276-279 then :68-88) -- a wasted draw whose value never reaches a density;
    it returns outcome 0 without consuming the stream (the oracle has no such draw)."""
    import venture.lite.dirichlet as dc
    assert K != V

    def sample(np_rng, alpha, counts):
        n = len(alpha)
        if n == V:
            return 0
        assert n == K
        return int(np_rng.tape.draw(4, [alpha[0]] + [counts[i] for i in range(n)]))
    dc.sample = sample


def lda_corpus(K, V, n_docs, n_per, seed):
    rng = np.random.default_rng(seed)
    theta = rng.dirichlet([0.5] * K, size=n_docs)
    phi = rng.dirichlet([0.5] * V, size=K)
    doc, word = [], []
    for d in range(n_docs):
        for _ in range(n_per):
            z = rng.choice(K, p=theta[d])
            doc.append(d)
            word.append(int(rng.choice(V, p=phi[z])))
    return doc, word


def chain_c4(seed=31, P=4, K=3, V=5, n_docs=4, n_per=3, n_mh=120):
    doc, word = lda_corpus(K, V, n_docs, n_per, 3)
    n_tok = len(doc)
    patch_dircat_sampler(K, V)
    out = {"seed": seed, "P": P, "K": K, "V": V, "n_docs": n_docs, "doc": doc, "word": word, "n_mh": n_mh,
           "a": 1.0, "b": 2.0, "steps": {"prior": 2, "tokens": 3, "mh_one": 40, "mh_each": 400}, "particles": []}
    for p in range(P):
        m, ad, aw = lda_model(K, V, seed, p, 2)
        rec = {"alpha_prior": [fl(m.value(ad)), fl(m.value(aw))], "incr": []}
        znodes = []
        pos = {}
        for i in range(n_tok):
            d = doc[i]
            pos[d] = pos.get(d, -1) + 1
            m.tape.begin(3, i)                     # z_i is draw i of step `tokens`
            did = m.observe(["word", d, pos[d]], vd.integer(word[i]))
            rec["incr"].append(fl(m.t.makeConsistent()))
            # (word d pos) -> mem request -> application of the lambda -> its body ((topic ((doc doc_id))))
            body = m.t.families[did]
            while not hasattr(body.operatorNode, "operandNodes"):
                esr = m.t.valueAt(body.requestNode).esrs[0]
                body = m.t.spFamilyAt(body.requestNode, esr.id)
            znodes.append(body.operatorNode.operandNodes[0])
        sites = [m.t.families[ad], m.t.families[aw]] + znodes
        m.tape.site_of = lambda key: [i for i, n in enumerate(sites) if n is key][0]

        def state():
            return {"alpha": [fl(m.value(ad)), fl(m.value(aw))],
                    "z": [int(m.t.valueAt(n).getNumber()) for n in znodes]}
        rec["after_tokens"] = state()
        rec["nodes_one"] = mh_steps(m, 40, n_mh)
        rec["after_one"] = state()
        out["particles"].append(rec)
    return out


# ------------------------------------------------------------------ C5 subsampled MH (austerity)
def chain_c5(seed=41, P=3, D=4, N=60, K=40, Nbatch=10, k0=2, eps=0.05):
    import ctypes
    import venture.lite.infer.subsampled_mh as ss
    X, y = blr_data(N, D, 50)
    out = {"seed": seed, "P": P, "D": D, "N": N, "K": K, "Nbatch": Nbatch, "k0": k0, "eps": eps, "data_seed": 50,
           "X": X.tolist(), "y": y.tolist(), "steps": {"prior": 0, "ssmh": 7}, "particles": []}
    tests = []
    orig = ss.sequentialTest

    def recording(*a):                      # reads sequentialTest's own return value; changes nothing
        r = orig(*a)
        tests.append(r)
        return r
    ss.sequentialTest = recording
    for p in range(P):
        m = Model()
        install(m, seed, p)
        m.tape.begin(0, 0)                  # the D prior normals: draws 2j, 2j+1 of step `prior`
        wd = [m.assume("w_%d" % j, ["normal", 0.0, 1.0]) for j in range(D)]
        m.assume("w", ["array"] + ["w_%d" % j for j in range(D)])
        tape = m.tape
        m.t.py_rng = __import__("random").Random(5)    # the flips evaluated by `observe` draw from py_rng; constrained next
        for i in range(N):
            m.observe(["log_odds_flip", ["vector_dot", "w", ["array"] + [float(v) for v in X[i]]]], vd.boolean(bool(y[i])))
        m.t.makeConsistent()
        m.t.py_rng = PyRng(tape)
        nodes = [m.t.families[d] for d in wd]
        tape.site_of = lambda key: [i for i, n in enumerate(nodes) if n is key][0]

        def permutation(items, tape=tape):   # subsampled_mh.py:103 -> the kernels' Feistel permutation of (seed, step)
            n = len(items)
            perm = np.empty(n, dtype=np.uint64)
            oracle.lib().orc_feistel_perm(ctypes.c_uint64(tape.seed), ctypes.c_uint32(tape.step), ctypes.c_uint64(n),
                                          perm.ctypes.data_as(ctypes.c_void_p))
            return [items[int(q)] for q in perm]
        m.t.np_rng.permutation = permutation
        rec = {"w_prior": [fl(m.value(d)) for d in wd]}
        del tests[:]
        rec["nodes"] = mh_steps(m, 7, K, op="subsampled_mh", extra=(Nbatch, k0, eps, False, 0, False))
        rec["accept"] = [bool(t[0]) for t in tests]
        rec["n"] = [float(t[1]) for t in tests]
        rec["tstat"] = [fl(t[2]) for t in tests]
        rec["w"] = [fl(m.value(d)) for d in wd]
        out["particles"].append(rec)
    ss.sequentialTest = orig
    return out


def chain_c4_gibbs(seed=37, P=4, K=3, V=5, n_docs=4, n_per=3, n_gibbs=60):
    """lda.vnt with z tagged into scope `z` (the default scope also holds the two gamma sites, which
    cannot be enumerated): (gibbs z one 1) x n_gibbs, then two ordered sweeps (gibbs z <block> 1)."""
    doc, word = lda_corpus(K, V, n_docs, n_per, 3)
    n_tok = len(doc)
    patch_dircat_sampler(K, V)
    out = {"seed": seed, "P": P, "K": K, "V": V, "n_docs": n_docs, "doc": doc, "word": word, "n_gibbs": n_gibbs,
           "a": 1.0, "b": 2.0, "steps": {"prior": 2, "tokens": 3, "gibbs_one": 40, "gibbs_each": 400}, "particles": []}
    for p in range(P):
        m, ad, aw = lda_model(K, V, seed, p, 2, tag_z=True)
        rec = {}
        blocks, pos = [], {}
        for i in range(n_tok):
            d = doc[i]
            pos[d] = pos.get(d, -1) + 1
            m.tape.begin(3, i)
            m.observe(["word", d, pos[d]], vd.integer(word[i]))
            m.t.makeConsistent()
            blocks.append(1000 * d + pos[d])
        znodes = []
        for b in blocks:
            pn = list(m.t.getNodesInBlock(vv.VentureSymbol("z"), vv.VentureNumber(b)))
            assert len(pn) == 1
            znodes.append(pn[0])
        m.tape.site_of = lambda key: blocks.index(int(key.getNumber()))

        def zs():
            return [int(m.t.valueAt(n).getNumber()) for n in znodes]
        rec["z_tokens"] = zs()
        rec["alpha"] = [fl(m.value(ad)), fl(m.value(aw))]
        rec["nodes_one"] = []
        for k in range(n_gibbs):
            m.tape.begin(40 + k, 2)
            rec["nodes_one"].append(m.t.primitive_infer(infer_exp("gibbs", "z", "one", 1)))
        rec["z_one"] = zs()
        rec["nodes_each"] = []
        for k in range(2 * n_tok):
            m.tape.begin(400 + k, 2)
            rec["nodes_each"].append(m.t.primitive_infer(infer_exp("gibbs", "z", blocks[k % n_tok], 1)))
        rec["z_each"] = zs()
        out["particles"].append(rec)
    return out


# ------------------------------------------------------------------ resample (trace_set.py:202-219)
def chain_resample(seed=5):
    """TraceSet._resample_traces draws each child's parent with sampleLogCategorical(log_weights, np_rng)
    (utils.py:170-180: max-shift, exp, normalise, np_rng.multinomial(1, ps).argmax()).  Here that loop runs
    on the reference's own functions with a stand-in np_rng whose multinomial inverts the CDF of the ps it is
    handed at a prescribed uniform U_k; the fixture records U and the parents.  New weights:
    log_domain_even_out (utils.py:152-157)."""
    from venture.lite.utils import log_domain_even_out, sampleLogCategorical
    rng = np.random.default_rng(seed)
    ninf = float("-inf")
    cases = {
        "random": (rng.normal(size=48) * 3).tolist(),
        "wide_range": (rng.normal(size=33) * 200).tolist(),
        "some_impossible": [ninf if i % 3 == 0 else float(v) for i, v in enumerate(rng.normal(size=40))],
        "all_impossible": [ninf] * 17,
        "ties": [0.0] * 25,
        "one_heavy": [0.0] * 10 + [30.0] + [0.0] * 10,
        "single": [-3.5],
    }

    class Inv(object):
        def __init__(self, us):
            self.us, self.k = us, 0

        def multinomial(self, n, ps):
            cum = np.cumsum(np.asarray(ps, dtype=np.float64))
            i = int(np.searchsorted(cum, self.us[self.k] * cum[-1], side="right"))
            self.k += 1
            out = np.zeros(len(ps), dtype=int)
            out[min(i, len(ps) - 1)] = 1
            return out
    out = {}
    for name, lw in cases.items():
        n_new = len(lw)
        us = rng.random(n_new).tolist()
        inv = Inv(us)
        parents = [int(sampleLogCategorical(lw, inv)) for _ in range(n_new)]    # trace_set.py:215-216
        out[name] = {"lw": [fl(x) for x in lw], "uniforms": us, "parents": parents,
                     "new_lw": [fl(x) for x in log_domain_even_out(lw, n_new)]}
    return out


CHAINS = [("c1", chain_c1), ("c2", chain_c2), ("c3", chain_c3), ("c4", chain_c4), ("c5", chain_c5), ("c4_gibbs", chain_c4_gibbs), ("resample", chain_resample)]


def main():
    only = sys.argv[1:]
    path = os.path.join(HERE, "lite_chains.json")
    fixture = {}
    if only and os.path.exists(path):
        with open(path) as f:
            fixture = json.load(f)
    for name, fn in CHAINS:
        if only and name not in only:
            continue
        fixture[name] = fn()
        print("ran", name)
    fixture["_source"] = ("reference Lite backend (backend/lite/*.py, py3 rendering by "
                          "oracle/lite_py3/build_lite_py3.py) run by tests/golden/make_lite_chains.py; "
                          "numpy %s" % np.__version__)
    with open(path, "w") as f:
        json.dump(fixture, f, indent=1, sort_keys=True)
    print("wrote", path)


if __name__ == "__main__":
    main()
