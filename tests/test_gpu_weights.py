"""GPU parity: logsumexp / normalised probs / resampling vs the oracle.
Reference semantics: backend/lite/utils.py:70-118,152-180; trace_set.py:202-235."""
import ctypes
import math

import numpy as np
import pytest

import venturecxx_b200 as vb
from venturecxx_b200 import _lib

pytestmark = pytest.mark.gpu


def gpu_lse(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    out = ctypes.c_double()
    _lib.call("vk_logsumexp_array", _lib.ptr(a) if a.size else None, a.size, 0, ctypes.byref(out))
    return out.value


def test_logsumexp_reference_vectors():
    # test/unit/test_explog.py:86-100
    inf, nan = float("inf"), float("nan")
    assert abs(gpu_lse(np.arange(1000.0)) - 999.4586751453871) / 999.4586751453871 < 1e-15
    assert gpu_lse([]) == -inf
    assert gpu_lse([-1000.0]) == -1000.0
    assert gpu_lse([-1000.0, -1000.0]) == -1000.0 + math.log(2.0)
    assert abs(gpu_lse([0.0, 0.0]) - math.log(2.0)) / math.log(2.0) < 1e-15
    assert gpu_lse([-inf, 1]) == 1
    assert gpu_lse([-inf, -inf]) == -inf
    assert gpu_lse([inf, inf]) == inf
    assert math.isnan(gpu_lse([-inf, inf]))
    assert math.isnan(gpu_lse([nan, inf]))
    assert math.isnan(gpu_lse([nan, -3]))


@pytest.mark.parametrize("n", [1, 31, 1000, 4097, 1 << 18])
def test_logsumexp_vs_oracle(orc, n):
    rng = np.random.default_rng(n)
    a = rng.normal(-50, 30, n)
    ref = orc.logsumexp(a)
    assert abs(gpu_lse(a) - ref) <= 1e-12 * abs(ref)


def make_ts(P, seed=7):
    return vb.B200TraceSet(vb.NormalNormalModel(), P, seed=seed)


@pytest.mark.parametrize("P", [1, 5, 1000, 2049, 1 << 16])
@pytest.mark.parametrize("mode", ["systematic", "multinomial"])
def test_resample_ancestors_bit_exact(orc, P, mode):
    rng = np.random.default_rng(P)
    ts = make_ts(P)
    lw = rng.normal(0, 3, P)
    if P > 4:
        lw[1] = -np.inf
    ts.set_particle_log_weights(lw)
    x_before = ts.read_site(0)
    step = ts.rng_step
    anc = ts.resample(P, mode=mode, return_ancestors=True)
    ref = orc.ancestors(lw, P, 1 if mode == "systematic" else 0, seed=7, step=step)
    assert np.array_equal(anc, ref)
    if mode == "systematic":
        assert np.all(np.diff(anc) >= 0)
    # ancestor state copied, weights evened out (utils.py:152-157)
    assert np.array_equal(ts.read_site(0), x_before[anc])
    lse = orc.logsumexp(lw)
    np.testing.assert_allclose(ts.log_weights, np.full(P, lse - math.log(P)), rtol=1e-12)
    if P > 4:
        assert 1 not in set(anc.tolist())


def test_resample_given_uniforms(orc):
    P = 1000
    rng = np.random.default_rng(3)
    lw = rng.normal(0, 2, P)
    ts = make_ts(P)
    ts.set_particle_log_weights(lw)
    anc = ts.resample(P, mode="systematic", uniforms=[0.25], return_ancestors=True)
    assert np.array_equal(anc, orc.ancestors(lw, P, 1, uniforms=[0.25]))
    ts.set_particle_log_weights(lw)
    u = rng.random(P)
    anc = ts.resample(P, mode="multinomial", uniforms=u, return_ancestors=True)
    assert np.array_equal(anc, orc.ancestors(lw, P, 0, uniforms=u))


@pytest.mark.parametrize("P", [1000, 1 << 16, (1 << 18) + 3])
def test_systematic_thresholds_on_cdf_steps(orc, P):
    # equal weights: child k's threshold (k + u) W / P sits EXACTLY on a CDF step when u = 0 -- the case
    # the kernels' double-precision shortcut must hand to the exact 128-bit test (k_weights.cu
    # first_child_at_least); also u just below 1 and weights that are exact powers of two apart
    rng = np.random.default_rng(11)
    for lw in [np.zeros(P), np.log(2.0) * rng.integers(0, 3, P).astype(np.float64)]:
        for u in [0.0, 0.5, 1.0 - 2.0 ** -32, 1.0 - 2.0 ** -53, 2.0 ** -32]:
            ts = make_ts(P)
            ts.set_particle_log_weights(lw)
            anc = ts.resample(P, mode="systematic", uniforms=[u], return_ancestors=True)
            assert np.array_equal(anc, orc.ancestors(lw, P, 1, uniforms=[u])), (P, u)


def test_resample_all_impossible_is_uniform(orc):
    # utils.py:93-99,176-180: all -inf => every parent equally likely
    P = 512
    ts = make_ts(P)
    ts.set_particle_log_weights(np.full(P, -np.inf))
    anc = ts.resample(P, mode="systematic", return_ancestors=True)
    assert np.array_equal(anc, np.arange(P))
    assert np.all(ts.log_weights == -np.inf)


def test_normalized_probs(orc):
    P = 777
    rng = np.random.default_rng(5)
    lw = rng.normal(0, 5, P)
    lw[3] = -np.inf
    ts = make_ts(P)
    ts.set_particle_log_weights(lw)
    np.testing.assert_allclose(ts.particle_normalized_probs(), orc.normalized_probs(lw), rtol=1e-12)
    ts.set_particle_log_weights(np.full(P, -np.inf))
    assert np.all(ts.particle_normalized_probs() == 0)
    assert ts.logsumexp() == -np.inf


@pytest.mark.parametrize("P_old,P_new", [(1, 1000), (1000, 333), (3000, 5000), (70000, 4096), (1500, 70000)])
@pytest.mark.parametrize("mode", ["systematic", "multinomial"])
def test_resample_changes_particle_count(orc, P_old, P_new, mode):
    """trace_set.py:202-219: (resample P_new) on P_old traces -- how `(resample 1000)` grows one particle into
    the particle set of the reference's programs (benchmarks/Makefile:6-9)."""
    rng = np.random.default_rng(P_old + P_new)
    ts = make_ts(P_old)
    lw = rng.normal(0, 2, P_old)
    ts.set_particle_log_weights(lw)
    x_before = ts.read_site(0)
    step = ts.rng_step
    anc = ts.resample(P_new, mode=mode, return_ancestors=True)
    ref = orc.ancestors(lw, P_new, 1 if mode == "systematic" else 0, seed=7, step=step)
    assert anc.shape == (P_new,) and np.array_equal(anc, ref)
    assert ts.n_particles == P_new
    assert np.array_equal(ts.read_site(0), x_before[anc])
    lse = orc.logsumexp(lw)
    got = ts.log_weights
    assert got.shape == (P_new,)
    np.testing.assert_allclose(got, np.full(P_new, lse - math.log(P_new)), rtol=1e-12)
    # the resized set keeps working: weights, a second (same-size) resample, prior resimulation
    ts.likelihood_weight()
    lw2 = ts.log_weights
    step = ts.rng_step
    anc2 = ts.resample(P_new, mode=mode, return_ancestors=True)
    assert np.array_equal(anc2, orc.ancestors(lw2, P_new, 1 if mode == "systematic" else 0, seed=7, step=step))


def test_resample_resizes_every_model_kind(orc):
    """state rows of every compiled model follow their particle through a size-changing resample"""
    from helpers import golden
    H = golden("hmm.json")
    hm = vb.B200TraceSet(vb.HMMModel(H["transition_matrix"], H["emission_matrix"], 1, 20), 700, seed=3)
    for d in H["data"][:5]:
        hm.observe(d); hm.incorporate()
    hist = np.stack([hm.read_site(t) for t in range(5)])
    anc = hm.resample(2500, mode="systematic", return_ancestors=True)
    assert np.array_equal(np.stack([hm.read_site(t) for t in range(5)]), hist[:, anc])
    hm.observe(H["data"][5]); hm.incorporate()
    assert hm.log_weights.shape == (2500,)
    X = np.random.default_rng(0).normal(size=(50, 3))
    y = (np.random.default_rng(1).random(50) < 0.5).astype(np.uint8)
    bl = vb.B200TraceSet(vb.LogisticRegressionModel(3, 50), 300, seed=4)
    bl.observe(X, y); bl.incorporate()
    w = np.stack([bl.read_site(j) for j in range(3)])
    ll = bl.log_likelihood_at("default", "all")
    anc = bl.resample(120, mode="multinomial", return_ancestors=True)
    assert np.array_equal(np.stack([bl.read_site(j) for j in range(3)]), w[:, anc])
    np.testing.assert_allclose(bl.log_likelihood_at("default", "all"), ll[anc], rtol=1e-12)
    nodes, acc = bl.resimulation_mh("default", "one", 10, return_accepts=True)
    assert nodes.shape == (120,) and acc.shape == (120,)
    st = bl.subsampled_mh("default", "one", 10, 2, 0.05, False, 0, False, 3)
    assert st.shape == (120, 3)


def test_resample_plus_infinite_weights_take_all_mass(orc):
    # max = +inf: the reference's sampleLogCategorical forms exp(inf - inf) and numpy raises; here the +inf
    # entries share the mass (the limit), so every child has an ancestor and nothing reads an unset index
    for P in (600, 5000):
        ts = make_ts(P)
        lw = np.random.default_rng(P).normal(0, 1, P)
        lw[[5, 77]] = np.inf
        ts.set_particle_log_weights(lw)
        step = ts.rng_step
        anc = ts.resample(P, mode="systematic", return_ancestors=True)
        assert set(anc.tolist()) == {5, 77}
        assert np.array_equal(anc, orc.ancestors(lw, P, 1, seed=7, step=step))
        assert np.all(ts.log_weights == np.inf)
