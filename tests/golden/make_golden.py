#!/usr/bin/env python
"""Generates tests/golden/*.json by EXECUTING THE REFERENCE'S OWN CODE for the
pure functions on the hot path, in this container (python 3.12, numpy/scipy as
installed).  Run from the repo root:  python tests/golden/make_golden.py

/root/reference is Python 2 and cannot be imported as a package; the functions
below are pure, so their source text is extracted verbatim from the reference
files (no reference source is copied into this repo -- only the outputs are
committed) and exec'd with small shims:
  * `venture.lite.{exception,mvnormal,typing}` stub modules for utils.py imports,
  * `xrange` -> `range` for sequentialTest,
  * base classes (`RandomPSP`, `DiscretePSP`, ...) replaced by `object`, and a
    MockArgs-like object (backend/lite/sp_use.py:30-62) for PSP methods.
The extracted line ranges are recorded in each fixture under "source".
Also extracts the literal golden tables the reference's tests/examples hold
(sequentialTest expectations, HMM matrices / data / filtered marginals).
"""
import json
import math
import os
import re
import sys
import types

import numpy as np
import scipy.special
import scipy.stats

REF = "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))


def read(rel):
    with open(os.path.join(REF, rel)) as f:
        return f.read()


def extract_block(rel, header):
    """Source text of the top-level (or nested) def/class starting at the line matching `header`,
    through the last following line that is blank or more indented."""
    lines = read(rel).split("\n")
    for i, ln in enumerate(lines):
        if re.match(header, ln):
            indent = len(ln) - len(ln.lstrip())
            j = i + 1
            while j < len(lines) and (lines[j].strip() == "" or len(lines[j]) - len(lines[j].lstrip()) > indent):
                j += 1
            block = lines[i:j]
            block = [b[indent:] for b in block]
            return "\n".join(block), "%s:%d-%d" % (rel, i + 1, j)
    raise KeyError((rel, header))


def jsonable(x):
    if isinstance(x, (list, tuple)):
        return [jsonable(v) for v in x]
    if isinstance(x, np.ndarray):
        return jsonable(x.tolist())
    if isinstance(x, (np.floating, float)):
        x = float(x)
        if math.isnan(x):
            return "nan"
        if math.isinf(x):
            return "inf" if x > 0 else "-inf"
        return x
    if isinstance(x, (np.integer, int, bool, np.bool_)):
        return int(x)
    return x


def dump(name, obj):
    with open(os.path.join(OUT, name), "w") as f:
        json.dump(obj, f, indent=1, sort_keys=True)
    print("wrote", name)


# ---------------------------------------------------------------- utils.py
def load_reference_utils():
    for name in ["venture", "venture.lite", "venture.lite.exception", "venture.lite.mvnormal", "venture.lite.typing"]:
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["venture.lite.exception"].VentureTypeError = type("VentureTypeError", (Exception,), {})
    import typing
    for n in ["Any", "Callable", "Type", "TypeVar"]:
        setattr(sys.modules["venture.lite.typing"], n, getattr(typing, n))
    sys.modules["venture.lite"].mvnormal = sys.modules["venture.lite.mvnormal"]
    sys.modules["venture"].lite = sys.modules["venture.lite"]
    mod = types.ModuleType("reference_lite_utils")
    exec(compile(read("backend/lite/utils.py"), "reference:backend/lite/utils.py", "exec"), mod.__dict__)
    return mod


def golden_utils():
    U = load_reference_utils()
    inf, nan = float("inf"), float("nan")
    rng = np.random.default_rng(20261104)
    lse_inputs = [[], [-1000.0], [-1000.0, -1000.0], [0.0, 0.0], [-inf, 1.0], [-inf, -inf], [inf, inf], [-inf, inf],
                  [nan, inf], [nan, -3.0], list(map(float, range(1000)))]
    for n in [3, 17, 256]:
        lse_inputs.append(rng.normal(-20, 15, n).tolist())
    xs = [0.0, -1.0, 1.0, 710.0, -710.0, 1000.0, -1000.0, 1e308, -1e308, -37.0, -36.999, -37.001, 36.0, 1e-9, -1e-9,
          inf, -inf] + rng.normal(0, 12, 40).tolist()
    cat_cases = []
    for n in [2, 5, 9]:
        ps = rng.random(n).tolist()
        ps[0] = 0.0
        for val in range(n):
            cat_cases.append({"ps": ps, "val": val})
    cat_cases.append({"ps": [0.0, 0.0, 0.0], "val": 1})       # normalizeList: all impossible -> uniform
    # objects array with a repeated object: mass of all equal positions is summed (utils.py:112-118)
    cat_obj = {"ps": [0.1, 0.2, 0.3, 0.4], "os": [1, 2, 1, 3], "vals": [1, 2, 3, 4]}
    dir_cases = []
    for n in [2, 4, 16]:
        alpha = (rng.random(n) * 3 + 0.2).tolist()
        theta = rng.dirichlet(alpha).tolist()
        dir_cases.append({"alpha": alpha, "theta": theta})
    lw_cases = [rng.normal(0, 5, 7).tolist(), [-inf, -inf, -inf], [-inf, 0.0, -2.0], [3.0]]
    out = {
        "source": "backend/lite/utils.py (whole module exec'd with stub imports)",
        "logsumexp": [{"in": jsonable(a), "out": jsonable(U.logsumexp(a))} for a in lse_inputs],
        "log_logistic": [{"x": jsonable(x), "out": jsonable(float(U.log_logistic(x)))} for x in xs],
        "logistic": [{"x": jsonable(x), "out": jsonable(float(U.logistic(x)))} for x in xs],
        "logit": [{"x": p, "out": jsonable(float(U.logit(p)))} for p in [0.5, 0.1, 0.9, 1e-9, 1 - 1e-9]],
        "logDensityCategorical": [dict(c, out=jsonable(U.logDensityCategorical(c["val"], c["ps"]))) for c in cat_cases],
        "logDensityCategorical_objects": dict(cat_obj, out=[jsonable(U.logDensityCategorical(v, cat_obj["ps"], cat_obj["os"]))
                                                             for v in cat_obj["vals"]]),
        "logDensityLogCategorical": [dict(c, out=jsonable(U.logDensityLogCategorical(
            c["val"], [U.log(p) for p in c["ps"]]))) for c in cat_cases[:-1]],
        "logDensityDirichlet": [dict(c, out=jsonable(float(U.logDensityDirichlet(c["theta"], c["alpha"])))) for c in dir_cases],
        "log_domain_even_out": [{"in": jsonable(a), "n": 5, "out": jsonable(U.log_domain_even_out(a, 5))} for a in lw_cases],
        "logWeightsToNormalizedDirect": [{"in": jsonable(a), "out": jsonable(U.logWeightsToNormalizedDirect(a))} for a in lw_cases],
        "normalizeList": [{"in": a, "out": jsonable(U.normalizeList(a))} for a in [[1.0, 3.0], [0.0, 0.0, 0.0, 0.0], [2.0]]],
    }
    # sampleLogCategorical frequency check input: the reference's multinomial resampler core (trace_set.py:209-219)
    lw = [-1.0, -2.5, 0.3, -inf, -0.2]
    np_rng = np.random.RandomState(7)
    draws = [U.sampleLogCategorical(lw, np_rng) for _ in range(20000)]
    out["sampleLogCategorical"] = {"logs": jsonable(lw), "n": 20000, "counts": np.bincount(draws, minlength=5).tolist(),
                                   "seed": "np.random.RandomState(7)"}
    dump("lite_utils.json", out)


# ------------------------------------------------------- sequentialTest
def golden_sequential_test():
    src, where = extract_block("backend/lite/infer/subsampled_mh.py", r"^def sequentialTest")
    ns = {"np": np, "stats": scipy.stats, "xrange": range, "math": math}
    exec(src, ns)
    sequentialTest = ns["sequentialTest"]
    tsrc = read("test/inference_language/test_subsampledmh.py")
    ys = eval(re.search(r"ys = (\[\[.*?\]\])", tsrc, re.S).group(1))
    accepts = [bool(v) for v in eval(re.search(r"accepts_expect = map\(bool, (\[.*?\])\)", tsrc).group(1))]
    ns_expect = eval(re.search(r"ns_expect = (\[.*?\])", tsrc).group(1))
    tstats = eval(re.search(r"tstats_expect = (\[.*?\])", tsrc).group(1))
    cases = []
    # the reference test's own table (test_subsampledmh.py:32-60)
    for y, a, n, t in zip(ys, accepts, ns_expect, tstats):
        acc, nn, tt = sequentialTest(0.2, 3, 2, 20, 0.1, lambda j: y[j])
        assert acc == a and nn == n and abs(tt - t) < 1e-5
        cases.append({"mu_0": 0.2, "k0": 3, "Nbatch": 2, "N": 20, "epsilon": 0.1, "dllh": y, "accept": int(acc),
                      "n": float(nn), "tstat": jsonable(tt), "tstat_expect_reference_test": t})
    # extra cases run through the reference function: other (k0, Nbatch, eps), constant sequences, n == N
    rng = np.random.default_rng(5)
    for (mu0, k0, nb, N, eps) in [(0.0, 1, 3, 30, 0.05), (-0.1, 2, 5, 23, 0.01), (0.3, 4, 1, 12, 0.2), (0.0, 3, 4, 16, 0.0)]:
        for _ in range(4):
            y = rng.normal(0.05, 1.0, N).round(3).tolist()
            acc, nn, tt = sequentialTest(mu0, k0, nb, N, eps, lambda j: y[j])
            cases.append({"mu_0": mu0, "k0": k0, "Nbatch": nb, "N": N, "epsilon": eps, "dllh": y, "accept": int(acc),
                          "n": float(nn), "tstat": jsonable(tt)})
    for const in [0.5, -0.5]:
        y = [const] * 10
        acc, nn, tt = sequentialTest(0.0, 2, 2, 10, 0.1, lambda j: y[j])
        cases.append({"mu_0": 0.0, "k0": 2, "Nbatch": 2, "N": 10, "epsilon": 0.1, "dllh": y, "accept": int(acc),
                      "n": float(nn), "tstat": jsonable(tt)})
    tcdf = [{"t": t, "df": df, "out": float(scipy.stats.t.cdf(t, df))}
            for t in [-8.0, -2.5, -0.3, 0.0, 0.7, 1.96, 5.0, 30.0] for df in [1, 2, 5, 17, 100, 65535]]
    dump("sequential_test.json", {"source": where, "cases": cases, "t_cdf_scipy": tcdf})


# ------------------------------------------------------------------ PSPs
class MockArgs(object):
    def __init__(self, vals, aux=None):
        self._vals, self._aux = vals, aux

    def operandValues(self): return self._vals
    def spaux(self): return self._aux


def golden_psps():
    out = {"source": {}}
    rng = np.random.default_rng(11)
    # continuous.py NormalOutputPSP.logDensityNumeric
    src, where = extract_block("backend/lite/continuous.py", r"^class NormalOutputPSP")
    body, _ = extract_block("backend/lite/continuous.py", r"^  def logDensityNumeric\(self,x,params\)")
    ns = {"math": math, "HALF_LOG2PI": 0.5 * math.log(2 * math.pi)}
    exec(body, ns)
    out["source"]["normal"] = where
    cases = [(float(x), float(mu), float(s)) for x, mu, s in zip(rng.normal(0, 5, 30), rng.normal(0, 3, 30), rng.random(30) * 4 + 0.01)]
    out["normal"] = [{"x": x, "mu": mu, "sigma": s, "out": ns["logDensityNumeric"](None, x, (mu, s))} for x, mu, s in cases]
    # gamma / beta delegate to scipy (continuous.py:1166-1167, 858-859): run those very calls
    gsrc, gwhere = extract_block("backend/lite/continuous.py", r"^  def logDensityNumeric\(self, x, alpha, beta\)")
    ns = {"scipy": __import__("scipy")}
    exec(gsrc, ns)
    out["source"]["gamma"] = gwhere
    gcases = [(float(x), float(a), float(b)) for x, a, b in zip(rng.gamma(2, 2, 30), rng.random(30) * 5 + 0.05, rng.random(30) * 4 + 0.05)]
    gcases += [(0.0, 1.0, 2.0), (0.0, 2.0, 2.0), (0.0, 0.5, 2.0), (-1.0, 2.0, 1.0), (1e-300, 0.3, 1.0), (50.0, 10.0, 10.0)]
    out["gamma"] = [{"x": x, "shape": a, "rate": b, "out": jsonable(float(ns["logDensityNumeric"](None, x, a, b)))} for x, a, b in gcases]
    bsrc, bwhere = extract_block("backend/lite/continuous.py", r"^  def logDensityNumeric\(self, x, params\):\n?")
    # several classes share that header; take BetaOutputPSP's (the one calling scipy.stats.beta.logpdf)
    text = read("backend/lite/continuous.py")
    i = text.index("class BetaOutputPSP")
    j = text.index("def logDensityNumeric", i)
    k = text.index("\n\n", j)
    bsrc = "\n".join(l[2:] for l in text[j - 2:k].split("\n"))
    exec(bsrc, ns)
    out["source"]["beta"] = "backend/lite/continuous.py BetaOutputPSP.logDensityNumeric (scipy.stats.beta.logpdf)"
    bcases = [(float(x), float(a), float(b)) for x, a, b in zip(rng.random(30), rng.random(30) * 5 + 0.05, rng.random(30) * 5 + 0.05)]
    bcases += [(0.5, 1.0, 1.0), (0.0, 1.0, 3.0), (1.0, 3.0, 1.0), (1.5, 2.0, 2.0), (1e-12, 0.5, 0.5), (0.25, 9.0, 0.1)]
    out["beta"] = [{"x": x, "a": a, "b": b, "out": jsonable(float(ns["logDensityNumeric"](None, x, (a, b))))} for x, a, b in bcases]

    # discrete.py flip family logDensity, executed with the reference's utils
    U = load_reference_utils()
    ns = {"log": U.log, "log1p": U.log1p, "log_logistic": U.log_logistic, "math": math, "DiscretePSP": object,
          "logit": U.logit}
    for cls in ["BernoulliOutputPSP", "LogBernoulliOutputPSP", "LogOddsBernoulliOutputPSP"]:
        text = read("backend/lite/discrete.py")
        i = text.index("class %s" % cls)
        j = text.index("  def logDensity", i)
        k = text.index("\n\n", j)
        fsrc = "\n".join(l[2:] for l in text[j:k].split("\n"))
        exec(fsrc, ns)
        fn = ns["logDensity"]
        if cls == "BernoulliOutputPSP":
            ps = [0.0, 1.0, 0.5, 0.25, 1e-12, 1 - 1e-12]
            out["flip"] = [{"p": p, "val": v, "out": jsonable(fn(None, bool(v), MockArgs([p])))} for p in ps for v in (0, 1)]
        elif cls == "LogBernoulliOutputPSP":
            lps = [-1e-9, -0.1, -1.0, -30.0, -800.0]
            out["log_flip"] = [{"logp": p, "val": v, "out": jsonable(fn(None, bool(v), MockArgs([p])))} for p in lps for v in (0, 1)]
        else:
            xs = [0.0, 1.0, -1.0, 40.0, -40.0, 710.0, -710.0, 3.3]
            out["log_odds_flip"] = [{"x": x, "val": v, "out": jsonable(float(fn(None, bool(v), MockArgs([x]))))} for x in xs for v in (0, 1)]
    out["source"]["flip family"] = "backend/lite/discrete.py Bernoulli/LogBernoulli/LogOddsBernoulli logDensity"

    # dirichlet.py CDirCatOutputPSP with range_tree.Node counts: logDensity + logDensityOfData
    rt = types.ModuleType("range_tree")
    exec(compile(read("backend/lite/range_tree.py"), "reference:range_tree.py", "exec"), rt.__dict__)
    text = read("backend/lite/dirichlet.py")
    i = text.index("class CDirCatOutputPSP")
    k = min(text.index("\nclass ", i + 10), text.index("\nregisterBuiltinSP", i))
    csrc = text[i:k]
    aux_i = text.index("class DirCatSPAux")
    aux_k = text.index("\nclass ", aux_i + 10)
    ns = {"RandomPSP": object, "SPAux": object, "Node": rt.Node, "sample": rt.sample, "math": math, "scipy": __import__("scipy"),
          "copy": __import__("copy")}
    exec(text[aux_i:aux_k], ns)
    exec(csrc, ns)
    out["source"]["dir_cat"] = "backend/lite/dirichlet.py DirCatSPAux + CDirCatOutputPSP, backend/lite/range_tree.py"
    dc = []
    for n, alpha in [(2, 0.5), (4, 1.0), (16, 0.37), (5, 2.5)]:
        counts = rng.integers(0, 6, n).tolist()
        psp = ns["CDirCatOutputPSP"]([alpha] * n, list(range(n)))
        aux = ns["DirCatSPAux"](counts=rt.Node(counts))
        dc.append({"alpha": alpha, "n": n, "counts": counts,
                   "logDensity": [psp.logDensity(v, MockArgs([], aux)) for v in range(n)],
                   "logDensityOfData": float(psp.logDensityOfData(aux))})
    out["sym_dir_cat"] = dc
    # range_tree.sample frequencies ~ (alpha + counts)
    np_rng = np.random.RandomState(3)
    counts = [3, 0, 1, 5, 0]
    tree_a, tree_c = rt.Node([0.5] * 5), rt.Node(counts)
    draws = [rt.sample(np_rng, tree_a, tree_c) for _ in range(20000)]
    out["range_tree_sample"] = {"alpha": 0.5, "counts": counts, "n": 20000, "freq": np.bincount(draws, minlength=5).tolist()}

    # discrete.py CBetaBernoulliOutputPSP logDensity / logDensityOfData
    text = read("backend/lite/discrete.py")
    i = text.index("class CBetaBernoulliOutputPSP")
    k = text.index("\n\nregisterBuiltinSP", i)
    ns = {"DiscretePSP": object, "math": math, "scipy": __import__("scipy")}
    exec(text[i:k], ns)

    class Aux(object):
        def __init__(self, y, n): self.yes, self.no = y, n
        def cts(self): return [self.yes, self.no]
    bb = []
    for a, b, y, n in [(1.0, 1.0, 100, 0), (1.0, 1.0, 0, 0), (0.5, 2.0, 3, 7), (3.0, 0.2, 12, 1)]:
        psp = ns["CBetaBernoulliOutputPSP"](a, b)
        aux = Aux(y, n)
        bb.append({"a": a, "b": b, "yes": y, "no": n, "logDensity_true": psp.logDensity(True, MockArgs([], aux)),
                   "logDensity_false": psp.logDensity(False, MockArgs([], aux)),
                   "logDensityOfData": float(psp.logDensityOfData(aux))})
    out["beta_bernoulli"] = bb
    out["source"]["beta_bernoulli"] = "backend/lite/discrete.py CBetaBernoulliOutputPSP"
    # test/conformance/test_get_global_logscore.py:32-50 known answers
    out["global_logscore"] = {"flip_half_x100": -100 * math.log(2), "beta_bernoulli_1_1_100_true": -math.log(101)}
    dump("lite_psps.json", out)


# ------------------------------------------------------------------- HMM
def golden_hmm():
    vnt = read("examples/hmm.vnt")
    t = 1.0 / 3.0  # hmm.vnt:32 (let ((t .33333333333333333)) ...)
    tm = re.search(r"\[assume transition_matrix(.*?)\]\n\n", vnt, re.S).group(1)
    rows = re.findall(r"\(simplex ([^)]*)\)", tm)
    trans = [[t if tok == "t" else float(tok) for tok in r.split()] for r in rows]
    em = re.search(r"\[assume emission_matrix(.*?)\]\n\n", vnt, re.S).group(1)
    emit = [[float(tok) for tok in r.split()] for r in re.findall(r"\(simplex ([^)]*)\)", em)]
    data = [int(v) for v in re.search(r"\[define data \(array ([^)]*)\)\]", vnt).group(1).split()]
    plug = read("examples/hmm_plugin.py")

    def table(name):
        m = re.search(name + r" = \[(.*?)\n\]", plug, re.S)
        txt = re.sub(r"#.*", "", m.group(1))
        return eval("[" + txt + "]")
    out = {"source": "examples/hmm.vnt:31-66, examples/hmm_plugin.py:130-205",
           "transition_matrix": trans, "emission_matrix": emit, "data": data, "init_state": 1,
           "filtered_marginals": table("filtered_marginals"), "smoothed_marginals": table("smoothed_marginals"),
           "maps": table("maps")}
    assert len(trans) == 6 and len(emit) == 5 and len(data) == 20 and len(out["filtered_marginals"]) == 20
    dump("hmm.json", out)


if __name__ == "__main__":
    golden_utils()
    golden_sequential_test()
    golden_psps()
    golden_hmm()
