#!/usr/bin/env python
"""Multi-GPU parity check over NCCL (one process per GPU):

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
      --master-port 29511 tools/check_multi_gpu.py

1. HMM particle filter (C3) with the particle set sharded over the ranks: after every
   sharded resample, this rank's histories must equal the matching slice of a single-GPU
   run over all particles (Philox streams are keyed by GLOBAL particle index and the
   fixed-point CDF makes ancestors independent of the GPU count) -- bit exact.
2. subsampled_mh (C5 shape, small N) with observations sharded and particles replicated, over NCCL
   (round-level vk_ssmh_* calls) and through the peer mailbox inside the persistent kernel: weights
   must equal the single-GPU run to 1e-12, accept counts and sections visited exactly.
Prints one JSON line on rank 0.  Timings at the BASELINE sizes: bench.py --gpus N (configs.c3, configs.c5).
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import venturecxx_b200 as vb  # noqa: E402
from venturecxx_b200.distributed import ShardedTraceSet  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    H = json.load(open(os.path.join(ROOT, "tests", "golden", "hmm.json")))
    out = {"world": world}

    # ---- 1. sharded HMM filter vs single GPU -------------------------------------
    for mode in ("systematic", "multinomial"):
        P_local = 40000 + 17
        P = world * P_local
        T = 8
        model = vb.HMMModel(H["transition_matrix"], H["emission_matrix"], 1, 20)
        sh = ShardedTraceSet(model, P_local, seed=77, device=local)
        ref = vb.B200TraceSet(model, P, seed=77, device=local)
        t_sh = 0.0
        for t, d in enumerate(H["data"][:T]):
            sh.observe(d); ref.observe(d)
            sh.incorporate(); ref.incorporate()
            assert sh.local.rng_step == ref.rng_step
            torch.cuda.synchronize(); dist.barrier()
            t0 = time.perf_counter()
            lse = sh.resample(mode=mode)
            torch.cuda.synchronize()
            t_sh += time.perf_counter() - t0
            lse_ref = ref.logsumexp()
            ref.resample(P, mode=mode)
            assert abs(lse - lse_ref) <= 1e-12 * abs(lse_ref), (lse, lse_ref)
            for tt in range(t + 1):
                a = sh.read_site(tt)
                b = ref.read_site(tt)[rank * P_local:(rank + 1) * P_local]
                assert np.array_equal(a, b), ("history differs", mode, t, tt, rank)
            np.testing.assert_allclose(sh.log_weights, ref.log_weights[rank * P_local:(rank + 1) * P_local], rtol=1e-12)
        out["hmm_" + mode] = {"particles": P, "steps": T, "sharded_resample_ms_per_step": 1e3 * t_sh / T}
        sh.local.close(); ref.close()

    # ---- 1b. the same filter with the NVLink peer-memory resample (no collective, no host sync) -------
    for mode in ("systematic", "multinomial"):
        P_local = 40000 + 17
        P = world * P_local
        T = 8
        model = vb.HMMModel(H["transition_matrix"], H["emission_matrix"], 1, 20)
        sh = ShardedTraceSet(model, P_local, seed=78, device=local)
        sh.enable_peer_memory()
        ref = vb.B200TraceSet(model, P, seed=78, device=local)
        for t, d in enumerate(H["data"][:T]):
            sh.observe(d); ref.observe(d)
            sh.incorporate(); ref.incorporate()
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
        out["hmm_peer_" + mode] = "bit-identical to the single-GPU run"
        dist.barrier()
        sh.local.close(); ref.close()

    # ---- 2. observation-sharded subsampled MH vs single GPU ------------------------
    N, D, P = 40000, 16, 64
    X = np.random.default_rng(4).normal(size=(N, D))
    w_true = np.random.default_rng(5).normal(size=D)
    y = (np.random.default_rng(6).random(N) < 1.0 / (1.0 + np.exp(-X @ w_true))).astype(np.uint8)
    n_loc = N // world
    lo = rank * n_loc
    hi = N if rank == world - 1 else lo + n_loc
    model = vb.LogisticRegressionModel(D, N)
    sh = ShardedTraceSet(model, P, seed=5, device=local, replicate_particles=True)
    sh.observe_shard(X[lo:hi], y[lo:hi], lo, N)
    sh.local.incorporate()
    ref = vb.B200TraceSet(model, P, seed=5, device=local)
    ref.observe(X, y)
    ref.incorporate()
    args = ("default", "one", 2048, 3, 0.01, False, 0, False, 30)
    ref.rng_step = 100; sh.local.rng_step = 100
    torch.cuda.synchronize(); dist.barrier()
    t0 = time.perf_counter()
    st_sh = sh.subsampled_mh(*args)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    st_ref = ref.subsampled_mh(*args)
    w_sh = np.stack([sh.local.read_site(j) for j in range(D)])
    w_ref = np.stack([ref.read_site(j) for j in range(D)])
    np.testing.assert_allclose(w_sh, w_ref, rtol=1e-12)
    assert np.array_equal(st_sh[:, 0], st_ref[:, 0]), "accept counts differ"
    np.testing.assert_allclose(st_sh[:, 1], st_ref[:, 1], rtol=1e-12)
    out["subsampled_mh"] = {"observations": N, "particles": P, "transitions": 30, "seconds": dt,
                            "accepted_mean": float(st_sh[:, 0].mean())}
    # ---- 2b. the same through the peer mailbox inside the persistent kernel (no collective, no host loop) ----
    sh2 = ShardedTraceSet(model, P, seed=5, device=local, replicate_particles=True)
    sh2.observe_shard(X[lo:hi], y[lo:hi], lo, N)
    sh2.local.incorporate()
    sh2.enable_peer_memory()
    sh2.local.rng_step = 100
    torch.cuda.synchronize(); dist.barrier()
    t0 = time.perf_counter()
    st_p = sh2.subsampled_mh(*args)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    w_p = np.stack([sh2.local.read_site(j) for j in range(D)])
    np.testing.assert_allclose(w_p, w_ref, rtol=1e-12)
    assert np.array_equal(st_p[:, 0], st_ref[:, 0]), "accept counts differ (peer mailbox)"
    assert np.array_equal(st_p[:, 1], st_ref[:, 1]), "sections visited differ (peer mailbox)"
    out["subsampled_mh_peer_memory"] = {"observations": N, "particles": P, "transitions": 30, "seconds": dt,
                                        "decisions": "identical to the single-GPU run"}
    dist.barrier()
    sh2.local.close()
    # ---- 3. timing of the sharded filter / subsampled MH at the BASELINE sizes: bench.py --gpus N (configs.c3, configs.c5)
    dist.barrier()
    if rank == 0:
        out["status"] = "ok"
        print(json.dumps(out))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
