"""GPU, one device: the shard primitives behind the multi-GPU resample
(vk_shard_*, vk_pack_state, vk_unpack_state) with two shards living on the same
GPU and the exchange done by hand -- the result must equal one unsharded batch.
Also the round-level subsampled-MH primitives against vk_subsampled_mh."""
import ctypes
import math

import numpy as np
import pytest
import torch

import oracle
import venturecxx_b200 as vb
from venturecxx_b200 import _lib
from venturecxx_b200 import distributed as D
from helpers import golden
from test_gpu_blr import make_data, read_w

pytestmark = pytest.mark.gpu
H = golden("hmm.json")


def manual_sharded_resample(shards, seed, step, mode):
    """what distributed.sharded_resample does, with the collectives replaced by python lists"""
    world = len(shards)
    P_local = shards[0].P
    P_total = world * P_local
    gmax = max(s.weight_max() for s in shards)
    parts = [s.build_cdf(gmax, D.frac_bits_for(P_total)) for s in shards]
    masses = [p[0] for p in parts]
    lse = D.lse_from_parts(gmax, [p[1] for p in parts])
    total = sum(masses)
    rows, slots = [[] for _ in range(world)], [[] for _ in range(world)]
    for r, s in enumerate(shards):
        off = sum(masses[:r])
        if mode == "systematic":
            k_lo, k_hi = D.systematic_child_range(off, masses[r], total, P_total, D.global_uniform(seed, step))
            anc = s.ancestors(off, total, P_total, k_lo, k_hi - k_lo, 1, step)
            ks = torch.arange(k_lo, k_hi, device=anc.device)
        else:
            a = s.ancestors(off, total, P_total, 0, P_total, 0, step)
            ks = torch.nonzero(a >= 0).flatten()
            anc = a[ks]
        assert bool((anc >= 0).all())
        packed = s.pack(anc.to(torch.int32))
        dest = torch.div(ks, P_local, rounding_mode="floor")
        for d in range(world):
            sel = dest == d
            rows[d].append(packed[sel])
            slots[d].append((ks[sel] - d * P_local).to(torch.int32))
    for d, s in enumerate(shards):
        sl = torch.cat(slots[d])
        assert sorted(sl.tolist()) == list(range(P_local))
        s.unpack(sl.contiguous(), torch.cat(rows[d]).contiguous())
        s.fill_log_weights(lse - math.log(P_total))
    return lse


@pytest.mark.parametrize("mode", ["systematic", "multinomial"])
def test_two_shards_equal_one_batch_hmm(mode):
    P_local, world, seed = 1500, 2, 71
    P = P_local * world
    model = lambda: vb.HMMModel(H["transition_matrix"], H["emission_matrix"], 1, 20)
    full = vb.B200TraceSet(model(), P, seed=seed)
    parts = [vb.B200TraceSet(model(), P_local, seed=seed, particle_base=r * P_local) for r in range(world)]
    for d in H["data"][:6]:
        for ts in [full] + parts:
            ts.observe(d)
            ts.incorporate()
        # per-particle work is independent of the sharding (global-index streams)
        np.testing.assert_array_equal(np.concatenate([p.read_site(full.num_sites() - 1) for p in parts]),
                                      full.read_site(full.num_sites() - 1))
        np.testing.assert_allclose(np.concatenate([p.log_weights for p in parts]), full.log_weights, rtol=1e-13)
        step = full.rng_step
        full.resample(P, mode=mode)
        for p in parts:
            p.rng_step = step + 1
        lse = manual_sharded_resample([D.CudaShard(p) for p in parts], seed, step, mode)
        for t in range(full.num_sites()):
            np.testing.assert_array_equal(np.concatenate([p.read_site(t) for p in parts]), full.read_site(t))
        np.testing.assert_allclose(np.concatenate([p.log_weights for p in parts]), full.log_weights, rtol=1e-12)


def test_two_shards_equal_one_batch_blr_rows():
    # logistic regression: the packed row carries w (SoA), the 8N-byte logit row and the cached log-likelihood
    N, Dm, P_local, seed = 200, 5, 96, 73
    X, y = make_data(N, Dm, seed=3)
    full = vb.B200TraceSet(vb.LogisticRegressionModel(Dm, N), 2 * P_local, seed=seed)
    parts = [vb.B200TraceSet(vb.LogisticRegressionModel(Dm, N), P_local, seed=seed, particle_base=r * P_local) for r in range(2)]
    for ts in [full] + parts:
        ts.observe(X, y)
        ts.incorporate()
    step = full.rng_step
    full.resample(2 * P_local, mode="systematic")
    for p in parts:
        p.rng_step = step + 1
    manual_sharded_resample([D.CudaShard(p) for p in parts], seed, step, "systematic")
    for ts in [full] + parts:
        ts.resimulation_mh("default", "one", 6)     # uses the migrated logit rows
    np.testing.assert_array_equal(np.concatenate([read_w(p, Dm) for p in parts], axis=1), read_w(full, Dm))
    np.testing.assert_allclose(np.concatenate([p.log_likelihood_at("default", "all") for p in parts]),
                               full.log_likelihood_at("default", "all"), rtol=1e-13)


def test_observation_shards_sum_to_unsharded_subsampled_mh():
    # two observation shards on one GPU; the all-reduce of (sum dllh, sum dllh^2) done by hand
    N, Dm, P, seed = 900, 6, 48, 75
    X, y = make_data(N, Dm, seed=5)
    Nbatch, k0, eps, K = 64, 2, 0.05, 12
    full = vb.B200TraceSet(vb.LogisticRegressionModel(Dm, N), P, seed=seed)
    full.observe(X, y)
    full.incorporate()
    cut = 400
    parts = []
    for lo, hi in [(0, cut), (cut, N)]:
        ts = vb.B200TraceSet(vb.LogisticRegressionModel(Dm, N), P, seed=seed)   # particles replicated
        ts.observe(X[lo:hi], y[lo:hi])
        ts.incorporate()
        _lib.call("vk_blr_set_observation_shard", ts._h, lo, N)
        parts.append(ts)
    w0 = read_w(full, Dm)
    for p in parts:
        np.testing.assert_array_equal(read_w(p, Dm), w0)
    stats_full = full.subsampled_mh("default", "one", Nbatch, k0, eps, False, 0, False, K)
    for p in parts:
        _lib.call("vk_ssmh_begin", p._h)
    for t in range(K):
        for p in parts:
            _lib.call("vk_ssmh_propose", p._h, _lib.BLOCK_ONE, t, 0, 0.0)
        for q0 in range(0, N, Nbatch):
            q1 = min(q0 + Nbatch, N)
            views = []
            for p in parts:
                ptr = ctypes.c_void_p()
                _lib.call("vk_ssmh_round", p._h, q0, q1, ctypes.byref(ptr))
                views.append(D._wrap_device(torch, ptr.value, P * 2, 0))
            tot = views[0] + views[1]
            active = []
            for p, v in zip(parts, views):
                v.copy_(tot)
                torch.cuda.synchronize()
                n = ctypes.c_int32()
                _lib.call("vk_ssmh_decide", p._h, ctypes.c_void_p(v.data_ptr()), q1, k0, eps, ctypes.byref(n))
                active.append(n.value)
            assert active[0] == active[1]
            if active[0] == 0:
                break
        for p in parts:
            _lib.call("vk_ssmh_commit", p._h)
    for p in parts:
        np.testing.assert_allclose(read_w(p, Dm), read_w(full, Dm), rtol=1e-12)
        st = np.empty((P, 3))
        _lib.call("vk_ssmh_stats", p._h, K, _lib.ptr(st))
        np.testing.assert_array_equal(st[:, 0], stats_full[:, 0])
        np.testing.assert_allclose(st[:, 1], stats_full[:, 1], rtol=1e-12)


@pytest.mark.parametrize("mode", ["systematic", "multinomial"])
def test_peer_memory_resample_world_of_one_matches_vk_resample(mode):
    """csrc/p2p_resample.cuh with world = 1 (own mailbox, own buffers, no IPC import): the whole kernel
    sequence -- exchanges, offspring / search into the ancestor array, gather, buffer swap -- must reproduce
    vk_resample bit for bit.  (world > 1 over NVLink: tools/check_multi_gpu.py.)"""
    import ctypes
    from venturecxx_b200 import _lib
    from helpers import golden
    H = golden("hmm.json")
    P, T = 50001, 10
    model = vb.HMMModel(H["transition_matrix"], H["emission_matrix"], 1, 20)
    a = vb.B200TraceSet(model, P, seed=91)
    b = vb.B200TraceSet(model, P, seed=91)
    buf = (ctypes.c_ubyte * (64 * 16))()
    n = ctypes.c_int32()
    _lib.call("vk_p2p_export", a._h, buf, 16, ctypes.byref(n))
    assert n.value == 4 + 2 * 2     # mailbox, ancestors, ss mailbox, ss sequence words, two buffers for each of (cur, hist)
    _lib.call("vk_p2p_open", a._h, 1, 0, buf, n.value)
    m = _lib.RESAMPLE_SYSTEMATIC if mode == "systematic" else _lib.RESAMPLE_MULTINOMIAL
    for t, d in enumerate(H["data"][:T]):
        a.observe(d); b.observe(d)
        a.incorporate(); b.incorporate()
        lse_ref = b.logsumexp()
        _lib.call("vk_p2p_resample", a._h, m)
        b.resample(P, mode=mode)
        e, lse = ctypes.c_int32(), ctypes.c_double()
        _lib.call("vk_p2p_status", a._h, ctypes.byref(e), ctypes.byref(lse))
        assert e.value == 0
        assert abs(lse.value - lse_ref) <= 1e-12 * abs(lse_ref)
        assert a.rng_step == b.rng_step
        for tt in range(t + 1):
            assert np.array_equal(a.read_site(tt), b.read_site(tt)), (t, tt)
        np.testing.assert_allclose(a.log_weights, b.log_weights, rtol=1e-12)
    # all weights impossible -> uniform ancestors (utils.py:95-99), same as the single-GPU path
    a.set_particle_log_weights(np.full(P, -np.inf)); b.set_particle_log_weights(np.full(P, -np.inf))
    _lib.call("vk_p2p_resample", a._h, m)
    b.resample(P, mode=mode)
    assert np.array_equal(a.read_site(T - 1), b.read_site(T - 1))
    assert np.all(np.isneginf(a.log_weights))
