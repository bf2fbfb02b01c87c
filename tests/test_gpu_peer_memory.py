"""GPU: the peer-memory paths (csrc/p2p_common.cuh) with a world of TWO processes.

On the one-GPU test box both ranks open cuda:0 and reach each other's buffers through CUDA IPC exactly as two
GPUs of a node would (the driver time-slices the two contexts, so every exchange costs a context switch: slow,
but the protocol -- mailboxes, ancestor stores into the owner's array, peer gathers, the subsampled-MH mailbox --
is the real one).  On a multi-GPU box (gpurun --gpus 2, tools/check_multi_gpu.py) rank r uses cuda:r.

  E2  sharded HMM particle filter, vk_p2p_resample, both resampling modes: histories bit-identical to ONE
      handle holding all particles, weights equal.
  E3  observation-sharded subsampled_mh through the peer mailbox inside the persistent kernel: chains,
      accept counts and sections visited equal to ONE handle holding all observations.
"""
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

pytestmark = pytest.mark.gpu


def _device_for(rank):
    n = torch.cuda.device_count()
    return rank if n > rank else 0


def _hmm_worker(rank, world, port, out_q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    os.environ.setdefault("VK_P2P_TIMEOUT_S", "60")
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import venturecxx_b200 as vb
    from venturecxx_b200.distributed import ShardedTraceSet
    from helpers import golden
    H = golden("hmm.json")
    dev = _device_for(rank)
    torch.cuda.set_device(dev)
    res = {}
    for mode in ("systematic", "multinomial"):
        P_local, T = 3000 + 7, 6
        P = world * P_local
        model = vb.HMMModel(H["transition_matrix"], H["emission_matrix"], 1, 20)
        sh = ShardedTraceSet(model, P_local, seed=78, device=dev)
        sh.enable_peer_memory()
        ref = vb.B200TraceSet(model, P, seed=78, device=dev)
        for t, d in enumerate(H["data"][:T]):
            sh.observe(d); ref.observe(d)
            sh.incorporate(return_increments=False); ref.incorporate()
            assert sh.local.rng_step == ref.rng_step
            sh.resample(mode=mode)
            timed_out, lse = sh.peer_status()
            assert not timed_out, "a peer never arrived at an exchange"
            lse_ref = ref.logsumexp()
            ref.resample(P, mode=mode)
            assert abs(lse - lse_ref) <= 1e-12 * abs(lse_ref), (lse, lse_ref)
            dist.barrier()                      # peers finished reading before anyone inspects / moves on
            for tt in range(t + 1):
                a = sh.read_site(tt)
                b = ref.read_site(tt)[rank * P_local:(rank + 1) * P_local]
                assert np.array_equal(a, b), ("peer-memory history differs", mode, t, tt, rank)
            np.testing.assert_allclose(sh.log_weights, ref.log_weights[rank * P_local:(rank + 1) * P_local], rtol=1e-12)
        res[mode] = "bit-identical"
        dist.barrier()
        sh.local.close(); ref.close()
    out_q.put((rank, res))
    dist.barrier()
    dist.destroy_process_group()


def _ssmh_worker(rank, world, port, out_q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    os.environ.setdefault("VK_P2P_TIMEOUT_S", "60")
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import venturecxx_b200 as vb
    from venturecxx_b200.distributed import ShardedTraceSet
    dev = _device_for(rank)
    torch.cuda.set_device(dev)
    N, D, P = 6000, 8, 48
    X = np.random.default_rng(4).normal(size=(N, D))
    w_true = np.random.default_rng(5).normal(size=D)
    y = (np.random.default_rng(6).random(N) < 1.0 / (1.0 + np.exp(-X @ w_true))).astype(np.uint8)
    n_loc = N // world
    lo = rank * n_loc
    hi = N if rank == world - 1 else lo + n_loc
    model = vb.LogisticRegressionModel(D, N)
    sh = ShardedTraceSet(model, P, seed=5, device=dev, replicate_particles=True)
    sh.observe_shard(X[lo:hi], y[lo:hi], lo, N)
    sh.local.incorporate()
    sh.enable_peer_memory()
    ref = vb.B200TraceSet(model, P, seed=5, device=dev)
    ref.observe(X, y)
    ref.incorporate()
    args = ("default", "one", 256, 3, 0.01, False, 0, False, 12)
    ref.rng_step = 100; sh.local.rng_step = 100
    dist.barrier()
    st_sh = sh.subsampled_mh(*args)
    st_ref = ref.subsampled_mh(*args)
    w_sh = np.stack([sh.local.read_site(j) for j in range(D)])
    w_ref = np.stack([ref.read_site(j) for j in range(D)])
    np.testing.assert_allclose(w_sh, w_ref, rtol=1e-12)
    assert np.array_equal(st_sh[:, 0], st_ref[:, 0]), "accept counts differ"
    assert np.array_equal(st_sh[:, 1], st_ref[:, 1]), "sections visited differ"
    fin = np.isfinite(st_ref[:, 2])
    np.testing.assert_allclose(st_sh[fin, 2], st_ref[fin, 2], rtol=1e-9)
    assert st_ref[:, 1].mean() < N, "the sequential tests should stop early"
    # every rank holds the same particles afterwards (identical decisions everywhere)
    gathered = [None] * world
    dist.all_gather_object(gathered, w_sh.tobytes())
    assert all(g == gathered[0] for g in gathered)
    out_q.put((rank, {"accepted_mean": float(st_sh[:, 0].mean())}))
    dist.barrier()
    sh.local.close(); ref.close()
    dist.destroy_process_group()


def _small_state_worker(rank, world, port, out_q):
    """models whose whole state is particle-minor (normal-normal; a compiled program with three sites): the fused
    cooperative resample kernel (default) or, with VK_P2P_UNFUSED=1 in the environment, the separate kernels"""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    os.environ.setdefault("VK_P2P_TIMEOUT_S", "60")
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import venturecxx_b200 as vb
    from venturecxx_b200.compiler import CompiledModel
    from venturecxx_b200.distributed import ShardedTraceSet
    dev = _device_for(rank)
    torch.cuda.set_device(dev)
    P_local = 5000 + 3
    P = world * P_local
    prog = [("assume", "a", ["normal", 0.0, 2.0]), ("assume", "b", ["normal", "a", 1.0]), ("assume", "s", ["gamma", 2.0, 2.0])] + \
           [("observe", ["normal", ["add", "a", "b"], "s"], 0.3 * i) for i in range(5)]
    res = {}
    for name, make in (("normal", lambda: vb.NormalNormalModel(0.5, 2.0, 1.5, 8)), ("program", lambda: CompiledModel(prog))):
        model = make()
        sh = ShardedTraceSet(model, P_local, seed=91, device=dev)
        sh.enable_peer_memory()
        ref = vb.B200TraceSet(model, P, seed=91, device=dev)
        for it in range(4):
            if name == "normal":
                sh.observe(1.0 + it); ref.observe(1.0 + it)
                sh.incorporate(return_increments=False); ref.incorporate()
            else:
                sh.likelihood_weight(); ref.likelihood_weight()
            sh.resample(mode="systematic")
            timed_out, lse = sh.peer_status()
            assert not timed_out
            ref.resample(P, mode="systematic")
            dist.barrier()
            for site in range(ref.num_sites()):
                a = sh.read_site(site)
                b = ref.read_site(site)[rank * P_local:(rank + 1) * P_local]
                assert np.array_equal(a, b), (name, it, site, rank)
            np.testing.assert_allclose(sh.log_weights, ref.log_weights[rank * P_local:(rank + 1) * P_local], rtol=1e-12)
        res[name] = "bit-identical"
        dist.barrier()
        sh.local.close(); ref.close()
    out_q.put((rank, res))
    dist.barrier()
    dist.destroy_process_group()


def _unfused_hmm_worker(rank, world, port, out_q):
    os.environ["VK_P2P_UNFUSED"] = "1"          # read once per process, at the first sharded resample
    _hmm_worker(rank, world, port, out_q)


def _unfused_small_state_worker(rank, world, port, out_q):
    os.environ["VK_P2P_UNFUSED"] = "1"
    _small_state_worker(rank, world, port, out_q)


def _absent_peer_worker(rank, world, port, out_q):
    """rank 1 never calls resample: rank 0's kernel gives up after VK_P2P_TIMEOUT_S, reports it, and the particle set
    refuses further use (include/venture_b200.h, failure policy of vk_p2p_resample)"""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    os.environ["VK_P2P_TIMEOUT_S"] = "1"
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import venturecxx_b200 as vb
    from venturecxx_b200.distributed import ShardedTraceSet
    from helpers import golden
    H = golden("hmm.json")
    dev = _device_for(rank)
    torch.cuda.set_device(dev)
    sh = ShardedTraceSet(vb.HMMModel(H["transition_matrix"], H["emission_matrix"], 1, 20), 2000, seed=3, device=dev)
    sh.enable_peer_memory()
    res = "idle"
    if rank == 0:
        sh.observe(H["data"][0])
        sh.incorporate(return_increments=False)
        sh.resample(mode="systematic")
        timed_out, _ = sh.peer_status()
        raised = False
        try:
            sh.log_weights
        except vb.VentureException:
            raised = True
        res = (timed_out, raised)
    dist.barrier()
    out_q.put((rank, res))
    dist.barrier()
    sh.local.close()
    dist.destroy_process_group()


def _run(worker, world=2):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() * 7 + hash(worker.__name__)) % 2000
    procs = [ctx.Process(target=worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=600)
    alive = [p for p in procs if p.is_alive()]
    for p in alive:
        p.kill()
    assert not alive, "a rank hung"
    assert [p.exitcode for p in procs] == [0] * world
    return sorted([q.get(timeout=10) for _ in range(world)], key=lambda t: t[0])


def test_peer_memory_resample_two_ranks():
    res = _run(_hmm_worker)
    assert all(r[1] == {"systematic": "bit-identical", "multinomial": "bit-identical"} for r in res)


def test_peer_memory_resample_small_state_models():
    res = _run(_small_state_worker)
    assert all(r[1] == {"normal": "bit-identical", "program": "bit-identical"} for r in res)


def test_peer_memory_resample_separate_kernels():
    """VK_P2P_UNFUSED=1: tile sums, offspring and gather as three launches (also what multinomial mode, wide rows and
    very large shards use)"""
    res = _run(_unfused_hmm_worker)
    assert all(r[1] == {"systematic": "bit-identical", "multinomial": "bit-identical"} for r in res)
    res = _run(_unfused_small_state_worker)
    assert all(r[1] == {"normal": "bit-identical", "program": "bit-identical"} for r in res)


def test_peer_memory_subsampled_mh_two_ranks():
    res = _run(_ssmh_worker)
    assert res[0][1] == res[1][1]


def test_peer_memory_absent_peer_is_reported():
    res = _run(_absent_peer_worker)
    assert res[0][1] == (True, True) and res[1][1] == "idle"
