#!/usr/bin/env python
"""Generates tests/golden/lite_aaa.json by calling THE REFERENCE'S OWN PSP / LKernel classes (imported from
oracle/_ref, the py3 rendering of backend/lite) for the makers SURVEY 8(a) A10 / A12 lists next to the symmetric
collapsed one: make_dir_cat (asymmetric, collapsed), make_uc_dir_cat, make_uc_sym_dir_cat, make_uc_beta_bernoulli.

For each maker: the made SP's logDensity for every outcome, logDensityOfData / marginalLogDensityOfData on
sufficient statistics, the maker's own logDensity of a drawn parameter, and the AAA LKernel's Gibbs draw and
weight (UDirCatAAALKernel dirichlet.py:250-263, USymDirCatAAALKernel :395-407, UBetaBernoulliAAALKernel
discrete.py:502-510) with the oracle's samplers behind np_rng (draw i of Philox(seed, i, step), as the batched SP
entry points address them), so the kernels must reproduce the drawn parameters.

  python oracle/lite_py3/build_lite_py3.py && python tests/golden/make_aaa_golden.py
"""
import json
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
import venture.lite.dirichlet as dm  # noqa: E402
import venture.lite.discrete as dd  # noqa: E402
from venture.lite.range_tree import Node  # noqa: E402
from venture.lite.value import VentureInteger  # noqa: E402

SP_DIRICHLET, SP_BETA = 9, 3


class Args(object):
    """what a PSP / LKernel reads from its args (psp.py:320-383)"""

    def __init__(self, vals, aux=None, made_aux=None, np_rng=None):
        self._vals, self._aux, self._made, self._np = vals, aux, made_aux, np_rng

    def operandValues(self): return self._vals
    def spaux(self): return self._aux
    def madeSPAux(self): return self._made
    def np_prng(self): return self._np


class NpStandIn(object):
    """np_rng whose dirichlet / beta are the oracle's samplers on stream (seed, row, step)"""

    def __init__(self, seed, row, step):
        self.seed, self.row, self.step = seed, row, step

    def _one(self, sp, params, dim):
        # row `row` of a batch whose other rows do not matter: build a batch of row+1 rows
        batch = np.tile(np.asarray(params, dtype=np.float64), (self.row + 1, 1))
        return oracle.sp_simulate(sp, batch, self.row + 1, dim=dim, seed=self.seed, step=self.step)[self.row]

    def dirichlet(self, alpha):
        return self._one(SP_DIRICHLET, alpha, len(alpha))

    def beta(self, a, b):
        return float(self._one(SP_BETA, [a, b], 1))


def theta_of(record):
    return [float(x) for x in record.sp.outputPSP.psp.theta]


def main():
    rng = np.random.default_rng(5)
    out = {"seed": 77, "step": 3}
    seed, step = out["seed"], out["step"]
    cases = []
    for row, (n, alpha) in enumerate([(3, [0.5, 1.5, 2.0]), (5, [1.0] * 5), (8, rng.random(8) * 3 + 0.1), (2, [0.2, 4.0])]):
        alpha = [float(a) for a in alpha]
        counts = [int(c) for c in rng.integers(0, 7, n)]
        os_ = [VentureInteger(i) for i in range(n)]
        aux = dm.DirCatSPAux(counts=Node(counts))
        c = {"n": n, "alpha": alpha, "counts": counts, "row": row}
        psp = dm.CDirCatOutputPSP(alpha, os_)                                     # make_dir_cat
        c["collapsed_logDensity"] = [float(psp.logDensity(o, Args([], aux))) for o in os_]
        c["collapsed_logDensityOfData"] = float(psp.logDensityOfData(aux))
        maker = dm.MakerUDirCatOutputPSP()                                        # make_uc_dir_cat
        rec = maker.simulate(Args([alpha], np_rng=NpStandIn(seed, row, step)))
        c["maker_theta"] = theta_of(rec)
        c["maker_logDensity"] = float(maker.logDensity(rec, Args([alpha])))
        c["marginalLogDensityOfData"] = float(maker.marginalLogDensityOfData(aux, Args([alpha])))
        upsp = rec.sp.outputPSP.psp
        c["uncollapsed_logDensity"] = [float(upsp.logDensity(o, None)) for o in os_]
        k = maker.getAAALKernel()
        rec2 = k.simulate(None, Args([alpha], made_aux=aux, np_rng=NpStandIn(seed, row, step + 1)))
        c["aaa_theta"] = theta_of(rec2)
        c["aaa_weight"] = float(k.weight(None, rec2, Args([alpha], made_aux=aux)))
        cases.append(c)
    out["dir_cat"] = cases
    sym = []
    for row, (n, alpha) in enumerate([(4, 0.5), (16, 0.37), (3, 2.5)]):
        counts = [int(c) for c in rng.integers(0, 5, n)]
        aux = dm.DirCatSPAux(counts=Node(counts))
        maker = dm.MakerUSymDirCatOutputPSP()                                     # make_uc_sym_dir_cat
        rec = maker.simulate(Args([alpha, n], np_rng=NpStandIn(seed, row, step)))
        k = maker.getAAALKernel()
        rec2 = k.simulate(None, Args([alpha, n], made_aux=aux, np_rng=NpStandIn(seed, row, step + 1)))
        sym.append({"n": n, "alpha": alpha, "counts": counts, "row": row, "maker_theta": theta_of(rec),
                    "maker_logDensity": float(maker.logDensity(rec, Args([alpha, n]))),
                    "marginalLogDensityOfData": float(maker.marginalLogDensityOfData(aux, Args([alpha, n]))),
                    "aaa_theta": theta_of(rec2), "aaa_weight": float(k.weight(None, rec2, Args([alpha, n], made_aux=aux)))})
    out["sym_dir_cat"] = sym
    bb = []
    for row, (a, b, y, n) in enumerate([(1.0, 1.0, 100, 0), (0.5, 2.0, 3, 7), (3.0, 0.2, 12, 1), (2.0, 2.0, 0, 0)]):
        aux = dd.SuffBernoulliSPAux()
        aux.yes, aux.no = y, n
        maker = dd.MakerUBetaBernoulliOutputPSP()                                 # make_uc_beta_bernoulli
        rec = maker.simulate(Args([a, b], np_rng=NpStandIn(seed, row, step)))
        w = float(rec.sp.outputPSP.psp.weight)
        k = maker.getAAALKernel()
        rec2 = k.simulate(None, Args([a, b], made_aux=aux, np_rng=NpStandIn(seed, row, step + 1)))
        w2 = float(rec2.sp.outputPSP.psp.weight)
        spsp = rec2.sp.outputPSP.psp
        bb.append({"a": a, "b": b, "yes": y, "no": n, "row": row, "maker_weight": w,
                   "maker_logDensity": float(maker.logDensity(rec, Args([a, b]))),
                   "marginalLogDensityOfData": float(maker.marginalLogDensityOfData(aux, Args([a, b]))),
                   "aaa_weight_param": w2, "aaa_weight": float(k.weight(None, rec2, Args([a, b], made_aux=aux))),
                   "suff_logDensity_true": float(spsp.logDensity(True, None)), "suff_logDensity_false": float(spsp.logDensity(False, None)),
                   "suff_logDensityOfData": float(spsp.logDensityOfData(aux))})
    out["beta_bernoulli"] = bb
    out["_source"] = ("backend/lite/dirichlet.py (CDirCatOutputPSP, MakerUDirCatOutputPSP, UDirCatAAALKernel, UDirCatOutputPSP, "
                      "MakerUSymDirCatOutputPSP, USymDirCatAAALKernel), backend/lite/discrete.py (MakerUBetaBernoulliOutputPSP, "
                      "UBetaBernoulliAAALKernel, SuffBernoulliOutputPSP), backend/lite/lkernel.py PosteriorAAALKernel; run by "
                      "tests/golden/make_aaa_golden.py on the py3 rendering; numpy %s" % np.__version__)
    with open(os.path.join(HERE, "lite_aaa.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print("wrote lite_aaa.json")


if __name__ == "__main__":
    main()
