"""The oracle (oracle/oracle.c) against chains produced by the reference's own Lite backend.

tests/golden/lite_chains.json was written by tests/golden/make_lite_chains.py, which runs
backend/lite (Trace, constructScaffold, detachAndExtract, regenAndAttach, mixMH, MHOperator, ...)
with stand-in RNG objects that hand Lite the draws of Philox stream (seed, particle, step).  Here the
oracle replays the same (seed, particle, step) and must end in the same state with the same weights:
1e-12 relative for reals, exact for discrete state and for Lite's returned node counts."""
import numpy as np
import pytest

import oracle
from helpers import golden, unj

RTOL = 1e-12
CH = golden("lite_chains.json")


def col(recs, key):
    return np.array([unj(r[key]) for r in recs], dtype=np.float64)


# ------------------------------------------------------------------ C1
def test_c1_normal_normal_chain():
    c = CH["c1"]
    P, K, seed = c["P"], c["K"], c["seed"]
    mu0, s0, s = c["mu0"], c["s0"], c["s"]
    y = np.array(c["y"])
    recs = c["particles"]
    L = oracle.lib()
    x, lw = np.empty(P), np.empty(P)
    y2 = np.ascontiguousarray(y[:2])
    L.orc_nn_likelihood_weight(oracle._ptr(x), oracle._ptr(lw), oracle._ptr(y2), oracle.I64(2), oracle.I64(P),
                               oracle.I64(0), oracle.F64(mu0), oracle.F64(s0), oracle.F64(s), oracle.U64(seed),
                               oracle.U32(c["steps"]["likelihood_weight"]))
    np.testing.assert_allclose(x, col(recs, "x_lw"), rtol=RTOL)
    np.testing.assert_allclose(lw, col(recs, "lw"), rtol=RTOL)
    incr = np.empty(P)
    L.orc_nn_make_consistent(oracle._ptr(x), oracle._ptr(lw), oracle._ptr(incr), oracle._ptr(y), oracle.I64(2),
                             oracle.I64(3), oracle.I64(P), oracle.F64(s))
    np.testing.assert_allclose(incr, col(recs, "incr"), rtol=RTOL)
    acc = np.empty(P, dtype=np.int64)
    L.orc_nn_mh(oracle._ptr(x), oracle._ptr(y), oracle.I64(3), oracle.I64(P), oracle.I64(0), oracle.F64(mu0),
                oracle.F64(s0), oracle.F64(s), oracle.U64(seed), oracle.U32(c["steps"]["mh"]), oracle.I32(K),
                oracle._ptr(acc))
    np.testing.assert_allclose(x, col(recs, "x_mh"), rtol=RTOL)
    assert acc.sum() > 0 and (acc < K).any()          # the chain both accepted and rejected
    for joint, key in [(0, "ll_at"), (1, "lj_at")]:
        out = np.empty(P)
        L.orc_nn_log_likelihood_at(oracle._ptr(x), oracle._ptr(y), oracle.I64(3), oracle.I64(P), oracle.F64(mu0),
                                   oracle.F64(s0), oracle.F64(s), oracle.I32(joint), oracle._ptr(out))
        np.testing.assert_allclose(out, col(recs, key), rtol=RTOL)
    # transloop's return value (dispatch.py:109-120): |regenCounts| = x + one absorbing node per observation
    for r in recs:
        assert r["nodes"] == [1.0 + len(y)] * K
