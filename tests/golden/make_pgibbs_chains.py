#!/usr/bin/env python
"""Generates tests/golden/pgibbs_chains.json: `(pgibbs seq ordered Pi 1)` -- PGibbsOperator under mixMH,
backend/lite/infer/pgibbs.py:49-126, dispatch.py:197-212 -- run on THE REFERENCE'S LITE BACKEND for
examples/hmm.vnt with the stand-in RNGs of make_lite_chains.py (every draw is the oracle's sampler at the
matching position of Philox stream (seed, particle, step): draw 1 mixMH's accept uniform, draws 2.. the
conditional SMC's draws in call order: per inner particle a state for the first block, then per block and
inner particle an ancestor (sampleLogCategorical over the Pi new weights + the retained path's) and a state,
then the final index).

  python oracle/lite_py3/build_lite_py3.py && python tests/golden/make_pgibbs_chains.py
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_lite_chains import hmm_model, infer_exp, install  # noqa: E402
import venture.lite.value as vv  # noqa: E402
import venture.value.dicts as vd  # noqa: E402


def chain(seed=61, P=5, K=8, Pi=6, T=9):
    data = [3, 2, 3, 5, 5, 1, 2, 5, 2, 4, 4, 1][:T]
    out = {"seed": seed, "P": P, "K": K, "Pi": Pi, "T": T, "data": data, "steps": {"filter": 1, "pgibbs": 100}, "particles": []}
    for p in range(P):
        m, rows, erows = hmm_model()
        install(m, seed, p)
        for t in range(1, T + 1):
            m.tape.begin(1 + (t - 1), 0)
            m.observe(["emission", ["state", t]], vd.number(data[t - 1]))
            m.t.makeConsistent()
        nodes = []
        for t in range(2, T + 1):
            pn = list(m.t.getNodesInBlock(vv.VentureSymbol("seq"), vv.VentureNumber(t)))
            nodes.append(pn[0])

        def hist():
            return [1] + [int(m.t.valueAt(n).getNumber()) for n in nodes]
        rec = {"hist_start": hist(), "nodes": [], "hists": []}
        for k in range(K):
            m.tape.begin(100 + k, 2)
            rec["nodes"].append(m.t.primitive_infer(infer_exp("pgibbs", "seq", "ordered", Pi, 1)))
            rec["hists"].append(hist())
        rec["draws_used_last"] = m.tape.d
        out["particles"].append(rec)
    out["trans"], out["emit"] = rows, erows
    return out


def main():
    fx = {"hmm": chain(), "hmm_one_particle": chain(seed=67, P=3, K=5, Pi=1, T=6)}
    fx["_source"] = "reference Lite backend (py3 rendering), PGibbsOperator, run by tests/golden/make_pgibbs_chains.py"
    with open(os.path.join(HERE, "pgibbs_chains.json"), "w") as f:
        json.dump(fx, f, indent=1, sort_keys=True)
    print("wrote pgibbs_chains.json")


if __name__ == "__main__":
    main()
