"""CPU, world_size 2 over gloo: the multi-GPU resampling plumbing
(venturecxx_b200/distributed.py) with a CPU stand-in for the CUDA shard built
from the oracle.  The sharded result must equal the single-shard oracle
resample of the concatenated particle set, bit for bit."""
import math
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import oracle  # noqa: E402
from venturecxx_b200 import distributed as D  # noqa: E402


class CpuShard(object):
    """Shard protocol on numpy + the oracle (what CudaShard does through the C ABI)."""

    def __init__(self, lw, state):
        self.lw, self.state, self.P = lw.copy(), state.copy(), lw.size
        self.cdf = None

    def weight_max(self):
        v = self.lw[~np.isnan(self.lw)]
        return float(v.max()) if v.size else -math.inf

    def weight_stats(self):
        v = self.lw[~np.isnan(self.lw)]
        return (float(v.max()) if v.size else -math.inf, float(v.min()) if v.size else math.inf,
                bool(np.isnan(self.lw).any()))

    def build_cdf(self, gmax, frac_bits):
        self.cdf = np.empty(self.P, dtype=np.uint64)
        oracle.lib().orc_fixed_cdf(oracle._ptr(self.lw), oracle.I64(self.P), oracle.F64(gmax), oracle.I32(frac_bits),
                                   oracle._ptr(self.cdf))
        with np.errstate(invalid="ignore"):
            se = float(np.nansum(np.exp(self.lw - gmax))) if math.isfinite(gmax) else 0.0
        return int(self.cdf[-1]), se

    def ancestors(self, offset, total, n_new, k_begin, k_count, mode, step):
        out = np.empty(max(k_count, 1), dtype=np.int32)
        oracle.lib().orc_shard_ancestors(oracle._ptr(self.cdf), oracle.I64(self.P), oracle.U64(offset), oracle.U64(total),
                                         oracle.I64(n_new), oracle.I64(k_begin), oracle.I64(k_count), oracle.I32(mode),
                                         oracle.U64(self.seed), oracle.U32(step), oracle._ptr(out))
        return torch.from_numpy(out[:k_count])

    def pack(self, particles):
        return torch.from_numpy(np.ascontiguousarray(self.state[particles.numpy().astype(np.int64)]))

    def unpack(self, slots, packed):
        new = np.empty_like(self.state)
        new[slots.numpy().astype(np.int64)] = packed.numpy()
        self.state = new

    def fill_log_weights(self, v):
        self.lw = np.full(self.P, v)

    def incorporate(self):
        return None

    def tensor(self, values, dtype):
        return torch.tensor(values, dtype=dtype)


def _worker(rank, world, port, mode, lw_all, state_all, seed, step, out_q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    P_local = lw_all.size // world
    sl = slice(rank * P_local, (rank + 1) * P_local)
    shard = CpuShard(lw_all[sl], state_all[sl])
    shard.seed = seed
    lse = D.sharded_resample(shard, seed, step, mode)
    out_q.put((rank, lse, shard.state, shard.lw))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["systematic", "multinomial"])
@pytest.mark.parametrize("case", ["generic", "one_heavy", "all_impossible", "with_nan", "mixed_inf", "plus_inf"])
def test_sharded_resample_equals_single_shard(mode, case):
    world, P_local, seed, step = 2, 257, 99, 7
    P = world * P_local
    rng = np.random.default_rng(1)
    lw = rng.normal(0, 3, P)
    if case == "one_heavy":
        lw[:] = -50.0
        lw[300] = 0.0          # every child descends from a particle on rank 1
    elif case == "all_impossible":
        lw[:] = -np.inf
    elif case == "with_nan":          # NaN weights carry no mass; the logsumexp is NaN (utils.py:77-91)
        lw[[3, 260, 400]] = np.nan
    elif case == "mixed_inf":         # +inf on one rank, -inf on the other: logsumexp is NaN on every rank
        lw[5] = np.inf
        lw[300] = -np.inf
    elif case == "plus_inf":          # +inf entries take all the mass; logsumexp = +inf
        lw[[7, 333]] = np.inf
    state = rng.integers(0, 255, size=(P, 48), dtype=np.uint8)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() + hash((mode, case))) % 2000
    procs = [ctx.Process(target=_worker, args=(r, world, port, mode, lw, state, seed, step, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in range(world)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    anc = oracle.ancestors(lw, P, 1 if mode == "systematic" else 0, seed=seed, step=step)
    new_state = np.concatenate([r[2] for r in res])
    assert np.array_equal(new_state, state[anc])
    lse = oracle.logsumexp(lw)
    for r in res:
        if math.isnan(lse):
            assert math.isnan(r[1]) and np.all(np.isnan(r[3]))
        elif math.isinf(lse):
            assert r[1] == lse and np.all(r[3] == lse)
        else:
            assert abs(r[1] - lse) <= 1e-12 * abs(lse)
            np.testing.assert_allclose(r[3], lse - math.log(P), rtol=1e-12)


def test_systematic_child_range_matches_brute_force():
    rng = np.random.default_rng(0)
    for _ in range(200):
        n_new = int(rng.integers(1, 300))
        masses = [int(v) for v in rng.integers(0, 1 << 40, size=int(rng.integers(1, 6)))]
        if sum(masses) == 0:
            masses[0] = 5
        total = sum(masses)
        u = float(rng.random())
        U0 = min(int(u * 4294967296.0), 0xFFFFFFFF)
        thr = [(((k << 32) + U0) * total) // (n_new << 32) for k in range(n_new)]
        off = 0
        covered = 0
        for m in masses:
            lo, hi = D.systematic_child_range(off, m, total, n_new, u)
            brute = [k for k in range(n_new) if off <= thr[k] < off + m]
            assert list(range(lo, hi)) == brute
            covered += hi - lo
            off += m
        assert covered == n_new


def test_host_philox_matches_oracle_stream():
    for seed, step in [(0, 0), (1234567890123, 17), (2 ** 63 + 5, 4000000000)]:
        ref = oracle.uniforms(seed, 0xFFFFFFFFFFFFFFFF, step, 1)[0]
        assert D.global_uniform(seed, step) == ref
    assert D.frac_bits_for(1000) == 52 and D.frac_bits_for(1 << 20) == 42 == oracle.lib().orc_frac_bits(oracle.I64(1 << 20))
