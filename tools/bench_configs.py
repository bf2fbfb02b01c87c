#!/usr/bin/env python
"""The BASELINE.json configs other than the headline (C2, which bench.py times itself), at the sizes of
SURVEY 8(d) D2.  bench.py imports the functions below and attaches what they return to its JSON line as
`configs.{c1,c3,c4,c5}`; run directly they print one JSON line each on one GPU:

  python tools/bench_configs.py [c1 c3 c4 c5] [--quick]

Every function takes a `ctx` (world, rank, local device, torch stream shared with the library, a reducer that
returns the MAX over ranks of a list of floats) so the same code is the 1-GPU and the N-GPU measurement:
  c1  normal-normal, 1000 particles, (resample, likelihood_weight, resample): independent replicas per GPU
  c3  HMM particle filter, resample every step: particles SHARDED over the GPUs, vk_p2p_resample
      (strong: 2^20 particles in total; weak: 2^20 per GPU); bit-identity against one GPU holding everything
  c4  LDA, 4096 particles per GPU: independent replicas (resimulation_mh default one; gibbs z each)
  c5  subsampled_mh on 10^7 observations SHARDED over the GPUs, 64 replicated particles, partial sums through
      the peer mailbox inside the persistent kernel; decisions compared with one GPU holding all observations
Device time is taken with CUDA events on the stream the library launches on.  cpu_baseline legs (rank 0, one
GPU only) time the oracle port -- or, for c1, the reference's own C++ program -- on a bounded sample.
"""
import json
import math
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def peak_gbs():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class Ctx(object):
    def __init__(self, world=1, rank=0, local=0, stream=None):
        import torch
        self.torch, self.world, self.rank, self.local = torch, world, rank, local
        self.stream = stream if stream is not None else torch.cuda.Stream()

    def barrier(self):
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, values):
        t = self.torch.tensor(list(values), dtype=self.torch.float64, device="cuda")
        if self.world > 1:
            import torch.distributed as dist
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(v) for v in t.tolist()]

    def min_over_ranks(self, values):
        return [-v for v in self.max_over_ranks([-v for v in values])]

    def timed(self, fn, sync=None):
        """device milliseconds of fn() on the shared stream, max over ranks; barrier on both sides"""
        torch = self.torch
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.barrier()
        e0.record(self.stream)
        fn()
        e1.record(self.stream)
        if sync is not None:
            sync()
        self.barrier()
        return self.max_over_ranks([e0.elapsed_time(e1)])[0]


def roofline(bytes_moved, seconds, n_gpus, note=None):
    peak, src = peak_gbs()
    gbs = bytes_moved / seconds / 1e9
    out = {"bound": "hbm", "achieved": gbs, "peak": peak * n_gpus, "unit": "GB/s", "frac": gbs / (peak * n_gpus),
           "traffic": None, "peak_source": src + (" x %d GPUs" % n_gpus if n_gpus > 1 else "")}
    if note:
        out["note"] = note
    return out


# ---------------------------------------------------------------------------------------------- C1
def c1_cpu_baseline(P=1000, seconds_target=4.0):
    """The reference's own native program for this model, benchmarks/normal-normal-chain.cxx `importance`
    (P prior draws + weights + one log-categorical draw per repetition), compiled unmodified against a
    <random>-backed GSL shim by oracle/Makefile into oracle/_ref/normal_normal_chain."""
    exe = os.path.join(ROOT, "oracle", "_ref", "normal_normal_chain")
    if not os.path.exists(exe):
        return None
    reps = 2000
    t0 = time.perf_counter()
    subprocess.run([exe, "2", "importance", str(reps), str(P)], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, check=True)
    dt = time.perf_counter() - t0
    reps2 = int(max(reps, min(200000, reps * seconds_target / max(dt, 1e-6))))
    t0 = time.perf_counter()
    subprocess.run([exe, "2", "importance", str(reps2), str(P)], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, check=True)
    dt = time.perf_counter() - t0
    return {"value": reps2 * P / dt, "unit": "particle-iterations/s", "cores": 1, "kind": "reference",
            "sample": "%d repetitions of `importance` over %d particles (the reference's benchmarks/normal-normal-chain.cxx, "
                      "compiled unmodified with std::mt19937 behind a GSL shim; one repetition = %d prior draws, weights and "
                      "ONE categorical draw, i.e. less work than the GPU arm's two P-child resamples), %.1f s incl. process start"
                      % (reps2, P, P, dt)}


def c1(ctx, P=1000, reps=1000, cpu=False):
    """(do (resample P) (likelihood_weight) (resample P)) x reps, normal-normal (benchmarks/Makefile:6-9)"""
    import venturecxx_b200 as vb
    ts = vb.B200TraceSet(vb.NormalNormalModel(0.0, 1.0, 1.0), 1, seed=1 + ctx.rank, device=ctx.local)
    ts.use_stream(ctx.stream.cuda_stream)
    ts.observe([2.0])
    ts.incorporate()
    ts.resample(P, mode="systematic")          # the reference's program grows one particle into P this way

    def run(n):
        for _ in range(n):
            ts.resample(P, mode="systematic"); ts.likelihood_weight(); ts.resample(P, mode="systematic")
    run(20)
    ms = ctx.timed(lambda: run(reps), ts.synchronize)
    x = ts.read_site(0)
    ts.close()
    dt = ms * 1e-3
    # D4: likelihood_weight 2s; resample 2s + 8 + 2*state(8) per particle, twice
    b = reps * P * (16 + 2 * (16 + 8 + 16)) * ctx.world
    out = {"workload": "C1 normal-normal, %d particles per GPU (independent replicas), (resample, likelihood_weight, resample)" % P,
           "metric": "particle program iterations/sec", "value": reps * P * ctx.world / dt, "unit": "particle-iterations/s",
           "n_gpus": ctx.world, "scaling": "weak (replicas)", "ms_per_step": ms / reps, "us_per_iteration": 1e3 * ms / reps,
           "gpu_launches_per_iteration": 3, "posterior_mean": float(x.mean()), "posterior_std": float(x.std()),
           "roofline": roofline(b, dt, ctx.world, "latency bound: 1000 particles are 24 KB of state; three launches per iteration")}
    if cpu and ctx.rank == 0:
        out["cpu_baseline"] = c1_cpu_baseline(P)
    return out


# ---------------------------------------------------------------------------------------------- C3
def hmm_golden():
    return json.load(open(os.path.join(ROOT, "tests", "golden", "hmm.json")))


def c3_cpu_baseline(H, P=1 << 16):
    import oracle
    from helpers import OracleHMM
    cores = os.cpu_count() or 1
    o = OracleHMM(H["transition_matrix"], H["emission_matrix"], 1, P, 32, 5)
    t0 = time.perf_counter()
    for t, d in enumerate(H["data"]):
        o.step(d, 2 * t)
        o.resample(oracle.ancestors(o.lw, P, 1, seed=5, step=2 * t + 1))
    dt = time.perf_counter() - t0
    T = len(H["data"])
    return {"value": P * T / dt, "unit": "particle-steps/s", "cores": cores, "kind": "port",
            "sample": "%d particles x %d filter steps: oracle orc_hmm_step (OpenMP over particles, %d threads) + orc_ancestors "
                      "(one thread) + numpy row gather, %.2f s" % (P, T, cores, dt)}


def _c3_filter(ctx, H, P_local, sharded, seed, check):
    """one timed filter (20 steps: observe, incorporate, resample); returns (ms, tvd, identical or None)"""
    import venturecxx_b200 as vb
    from venturecxx_b200.distributed import ShardedTraceSet
    torch = ctx.torch
    model = vb.HMMModel(H["transition_matrix"], H["emission_matrix"], 1, 20)
    T = len(H["data"])

    def make(s):
        if sharded:
            ts = ShardedTraceSet(model, P_local, seed=s, device=ctx.local)
            ts.local.use_stream(ctx.stream.cuda_stream)
            ts.enable_peer_memory()
        else:
            ts = vb.B200TraceSet(model, P_local, seed=s, device=ctx.local, particle_base=0)
            ts.use_stream(ctx.stream.cuda_stream)
        return ts

    def run(ts):
        for d in H["data"]:
            ts.observe(d); ts.incorporate(return_increments=False)
            if sharded:
                ts.resample(mode="systematic")
            else:
                ts.resample(P_local, mode="systematic")

    def close(ts):
        ctx.barrier()
        (ts.local if sharded else ts).close()

    ts = make(seed)                      # warm-up: module load, allocations, peer mappings
    run(ts)
    (ts.local if sharded else ts).synchronize()
    close(ts)
    ts = make(seed + 1)
    host = {}

    def run_timed():
        t0 = time.perf_counter()
        run(ts)
        host["enqueue_s"] = time.perf_counter() - t0      # host time to enqueue the whole filter (no synchronisation inside)
    ms = ctx.timed(run_timed, (ts.local if sharded else ts).synchronize)
    _c3_filter.last_host_us_per_step = 1e6 * host["enqueue_s"] / len(H["data"])
    last = ts.read_site(T - 1).astype(np.int64)
    counts = np.bincount(last, minlength=6)[1:6].astype(np.float64)
    tot = ctx.torch.tensor(counts, dtype=torch.float64, device="cuda")
    if ctx.world > 1 and sharded:
        import torch.distributed as dist
        dist.all_reduce(tot)
    m = tot.cpu().numpy()
    tvd = 0.5 * float(np.abs(m / m.sum() - np.array(H["filtered_marginals"][T - 1])).sum())
    identical = None
    if check and sharded:
        # the same filter on ONE handle holding every particle (this rank replays it): histories must be equal
        ref = vb.B200TraceSet(model, P_local * ctx.world, seed=seed + 1, device=ctx.local)
        for d in H["data"]:
            ref.observe(d); ref.incorporate(return_increments=False); ref.resample(P_local * ctx.world, mode="systematic")
        ok = 1.0
        for t in (T - 1, T // 2, 1):
            a = ts.read_site(t)
            b = ref.read_site(t)[ctx.rank * P_local:(ctx.rank + 1) * P_local]
            if not np.array_equal(a, b):
                ok = 0.0
        if not np.allclose(ts.log_weights, ref.log_weights[ctx.rank * P_local:(ctx.rank + 1) * P_local], rtol=1e-12, atol=0):
            ok = 0.0
        ref.close()
        identical = bool(ctx.min_over_ranks([ok])[0] > 0.5)
    close(ts)
    return ms, tvd, identical


def c3(ctx, P_total=1 << 20, cpu=False, weak=True):
    """HMM (examples/hmm.vnt) particle filter, resample every step; particles sharded over the GPUs."""
    H = hmm_golden()
    T = len(H["data"])
    sharded = ctx.world > 1
    ms, tvd, ident = _c3_filter(ctx, H, P_total // ctx.world, sharded, 5, True)
    dt = ms * 1e-3
    out = {"workload": "C3 HMM particle filter (examples/hmm.vnt), %d particles in total over %d GPU(s), %d steps of observe / "
                       "incorporate / systematic resample%s" % (P_total, ctx.world, T, " through vk_p2p_resample (peer memory, no collective)" if sharded else ""),
           "metric": "particle-steps/sec", "value": P_total * T / dt, "unit": "particle-steps/s", "n_gpus": ctx.world,
           "scaling": "strong", "ms_per_step": ms / T, "us_per_step": 1e3 * ms / T, "gpu_launches_per_step": 4,
           "tvd_last_vs_exact_filter": tvd, "bit_identical_to_1gpu": ident,
           "host_enqueue_us_per_step": _c3_filter.last_host_us_per_step,
           "roofline": roofline(P_total * T * 208.0, dt, ctx.world, "SURVEY 8(d) D4: 208 B per particle-step; single-wave kernels, latency bound at this size")}
    if weak and sharded:
        P_big = 1 << 20
        ms2, tvd2, _ = _c3_filter(ctx, H, P_big, True, 7, False)
        out["weak"] = {"particles_per_gpu": P_big, "particles": P_big * ctx.world, "value": P_big * ctx.world * T / (ms2 * 1e-3),
                       "unit": "particle-steps/s", "scaling": "weak", "ms_per_step": ms2 / T, "us_per_step": 1e3 * ms2 / T,
                       "tvd_last_vs_exact_filter": tvd2,
                       "roofline": roofline(P_big * ctx.world * T * 208.0, ms2 * 1e-3, ctx.world)}
    if cpu and ctx.rank == 0:
        out["cpu_baseline"] = c3_cpu_baseline(H)
    return out


# ---------------------------------------------------------------------------------------------- C4
def lda_corpus(K, V, n_docs, n_per, seed=3):
    rng = np.random.default_rng(seed)
    theta = rng.dirichlet(np.full(K, 0.5), n_docs)
    phi = rng.dirichlet(np.full(V, 0.5), K)
    doc, word = [], []
    for d in range(n_docs):
        for k in rng.choice(K, n_per, p=theta[d]):
            doc.append(d); word.append(rng.choice(V, p=phi[k]))
    return np.array(doc, dtype=np.int32), np.array(word, dtype=np.int32)


def c4_cpu_baseline(K, V, n_docs, doc, word, seconds_target=6.0):
    from helpers import OracleLDA
    cores = os.cpu_count() or 1
    n_tok = len(doc)
    o = OracleLDA(K, V, n_docs, n_tok, 1.0, 2.0, cores, 7)
    o.observe(doc, word)
    o.seq(0, 0, n_tok, 0)                  # prior simulation + incorporate of every token
    n_sites = n_tok + 2
    t0 = time.perf_counter()
    o.mh(10, n_sites, -1)
    dt = time.perf_counter() - t0
    sweeps = int(max(1, min(50, seconds_target / max(dt, 1e-6))))
    t0 = time.perf_counter()
    o.mh(10 + n_sites, n_sites * sweeps, -1)
    dt = time.perf_counter() - t0
    return {"value": cores * n_sites * sweeps / dt, "unit": "transitions/s", "cores": cores, "kind": "port",
            "sample": "%d particles (one per core, OpenMP) x %d sweeps of %d sites, oracle orc_lda_mh, %.1f s" % (cores, sweeps, n_sites, dt)}


def c4(ctx, P=4096, sweeps=4, cpu=False):
    import venturecxx_b200 as vb
    K, V, n_docs, n_per = 16, 1024, 64, 64
    doc, word = lda_corpus(K, V, n_docs, n_per)
    n_tok = len(doc)
    ts = vb.B200TraceSet(vb.LDAModel(K, V, n_docs, n_tok), P, seed=7, device=ctx.local, particle_base=ctx.rank * P)
    ts.use_stream(ctx.stream.cuda_stream)
    ts.observe(doc, word)
    ts.incorporate(return_increments=False)
    n_sites = n_tok + 2
    from venturecxx_b200 import _lib
    mh = lambda k: _lib.call("vk_mh", ts._h, 0, _lib.BLOCK_ONE, k, None, None)
    gibbs = lambda k: _lib.call("vk_gibbs", ts._h, 1, _lib.BLOCK_EACH, k, None, None)
    mh(n_sites)
    ms = ctx.timed(lambda: [mh(n_sites) for _ in range(sweeps)], ts.synchronize)
    gibbs(1)
    ms_g = ctx.timed(lambda: gibbs(sweeps), ts.synchronize)
    ts.close()
    tr = float(P) * ctx.world * n_sites * sweeps
    trg = float(P) * ctx.world * n_tok * sweeps
    out = {"workload": "C4 LDA (examples/lda.vnt, collapsed sym_dir_cat counts), K=16, V=1024, 64 docs x 64 tokens, %d particles "
                       "per GPU (independent replicas), sweeps of (resimulation_mh default one %d)" % (P, n_sites),
           "metric": "particle-site MH transitions/sec", "value": tr / (ms * 1e-3), "unit": "transitions/s", "n_gpus": ctx.world,
           "scaling": "weak", "ms_per_step": ms / sweeps, "gpu_launches_per_step": 1,
           "roofline": roofline(tr * 100.0, ms * 1e-3, ctx.world, "SURVEY 8(d) D4: ~100 B per z-site transition; the kernel keeps a particle's "
                                "row in shared memory for a whole sweep, it is bound by the serial dependency chain per particle"),
           "gibbs": {"metric": "enumerative-Gibbs z-site transitions/sec (gibbs z each)", "value": trg / (ms_g * 1e-3),
                     "unit": "transitions/s", "ms_per_sweep": ms_g / sweeps,
                     "roofline": roofline(trg * (12.0 * K + 22.0), ms_g * 1e-3, ctx.world)}}
    if cpu and ctx.rank == 0:
        out["cpu_baseline"] = c4_cpu_baseline(K, V, n_docs, doc, word)
    return out


# ---------------------------------------------------------------------------------------------- C5
def c5_data(N, D):
    X = np.random.default_rng(4).normal(size=(N, D))
    w_true = np.random.default_rng(5).normal(size=D)
    y = (np.random.default_rng(6).random(N) < 1.0 / (1.0 + np.exp(-X @ w_true))).astype(np.uint8)
    return X, y


def c5_cpu_baseline(D, Nbatch, k0, eps, N=10 ** 6, seconds_target=8.0):
    import oracle
    cores = os.cpu_count() or 1
    X, y = c5_data(N, D)
    L = oracle.lib()

    def run(Ps, Ts):
        w = np.ascontiguousarray(np.random.default_rng(3).normal(size=(D, Ps)))
        st = np.empty((Ps, 3))
        t0 = time.perf_counter()
        L.orc_blr_subsampled_mh_lazy(oracle._ptr(w), oracle._ptr(X), oracle._ptr(y), oracle.I32(D), oracle.I64(N), oracle.I64(Ps),
                                     oracle.I64(0), oracle.F64(0.0), oracle.F64(1.0), oracle.U64(9), oracle.U32(0), oracle.I32(Ts),
                                     oracle.I32(-1), oracle.F64(float(Nbatch)), oracle.I32(k0), oracle.F64(eps), oracle.I32(0),
                                     oracle.F64(0.0), oracle._ptr(st))
        return time.perf_counter() - t0, st
    dt, _ = run(cores, 2)
    Ts = int(max(2, min(40, 2 * seconds_target / max(dt, 1e-6))))
    dt, st = run(cores, Ts)
    return {"value": cores * Ts / dt, "unit": "transitions/s", "cores": cores, "kind": "port",
            "sample": "%d particles (one per core, OpenMP) x %d transitions on N = %d observations (1/10 of the GPU arm's; the "
                      "sequential test visits a similar NUMBER of sections), oracle orc_blr_subsampled_mh_lazy (sections "
                      "evaluated minibatch by minibatch as Lite does), mean sections visited %.0f, %.1f s"
                      % (cores, Ts, N, float(st[:, 1].mean()), dt)}


def c5(ctx, N=10 ** 7, P=64, K=100, Nbatch=65536, k0=3, eps=0.01, cpu=False, check=True):
    """(subsampled_mh default one 65536 3 0.01 false 0 false K): observations sharded, particles replicated."""
    import venturecxx_b200 as vb
    from venturecxx_b200.distributed import ShardedTraceSet
    D = 16
    X, y = c5_data(N, D)
    sharded = ctx.world > 1
    model = vb.LogisticRegressionModel(D, N)
    n_loc = N // ctx.world
    lo = ctx.rank * n_loc
    hi = N if ctx.rank == ctx.world - 1 else lo + n_loc
    if sharded:
        sh = ShardedTraceSet(model, P, seed=9, device=ctx.local, replicate_particles=True)
        sh.local.use_stream(ctx.stream.cuda_stream)
        sh.observe_shard(X[lo:hi], y[lo:hi], lo, N)
        sh.local.incorporate(return_increments=False)
        sh.enable_peer_memory()
        ts = sh.local
    else:
        ts = vb.B200TraceSet(model, P, seed=9, device=ctx.local)
        ts.use_stream(ctx.stream.cuda_stream)
        ts.observe(X, y)
        ts.incorporate(return_increments=False)
    from venturecxx_b200 import _lib
    run = lambda k, out=None: _lib.call("vk_subsampled_mh", ts._h, 0, _lib.BLOCK_ONE, int(Nbatch), int(k0), float(eps), 0, 0.0, 0,
                                        int(k), None, _lib.ptr(out))
    ts.rng_step = 1000
    run(3)
    ts.synchronize()
    ts.rng_step = 2000
    stats = np.empty((P, 3))
    ms = ctx.timed(lambda: run(K), ts.synchronize)
    ts.rng_step = 2000 + K
    _lib.call("vk_ssmh_stats", ts._h, K, _lib.ptr(stats))
    w_end = np.stack([ts.read_site(j) for j in range(D)])
    identical = None
    if sharded and check:
        ok = 1.0
        if ctx.rank == 0:
            # one GPU holding ALL observations replays the same transitions: decisions must be the same
            ref = vb.B200TraceSet(model, P, seed=9, device=ctx.local)
            ref.observe(X, y)
            ref.incorporate(return_increments=False)
            ref.rng_step = 1000
            ref.subsampled_mh("default", "one", Nbatch, k0, eps, False, 0, False, 3)
            ref.rng_step = 2000
            st_ref = ref.subsampled_mh("default", "one", Nbatch, k0, eps, False, 0, False, K)
            w_ref = np.stack([ref.read_site(j) for j in range(D)])
            ref.close()
            if not (np.array_equal(st_ref[:, 0], stats[:, 0]) and np.array_equal(st_ref[:, 1], stats[:, 1])
                    and np.allclose(w_ref, w_end, rtol=1e-12, atol=0)):
                ok = 0.0
        identical = bool(ctx.min_over_ranks([ok])[0] > 0.5)
    ctx.barrier()
    ts.close()
    dt = ms * 1e-3
    visited = float(stats[:, 1].sum()) * K            # local sections over all particles and transitions
    evals = 2.0 * visited                              # old and new log density of every visited section
    out = {"workload": "C5 subsampled_mh logistic regression, N = %d observations sharded over %d GPU(s), D = 16, %d replicated "
                       "particles, (subsampled_mh default one %d %d %g false 0 false %d)%s"
                       % (N, ctx.world, P, Nbatch, k0, eps, K, ", partial sums through the peer mailbox inside one persistent kernel" if sharded else ", one persistent kernel per call"),
           "metric": "subsampled-MH particle transitions/sec", "value": P * K / dt, "unit": "transitions/s", "n_gpus": ctx.world,
           "scaling": "strong", "ms_per_step": ms / K, "obs_loglik_evals_per_sec": evals / dt,
           "mean_sections_per_transition": float(stats[:, 1].mean()), "accepted_per_particle": float(stats[:, 0].mean()),
           "gpu_launches_per_step": 1.0 / K, "decisions_identical_to_1gpu": identical,
           "roofline": roofline(visited * (8.0 * D + 1.0) / min(P, 32), dt, ctx.world,
                                "SURVEY 8(d) D4: (8D+1)/P_tile B per (particle, observation) evaluation, P_tile = 32; the path is "
                                "bound by the LATENCY of the sequential tests (mean %.1f minibatches, slowest of %d particles many more), not by bytes"
                                % (float(stats[:, 1].mean()) / Nbatch, P))}
    if cpu and ctx.rank == 0:
        out["cpu_baseline"] = c5_cpu_baseline(D, Nbatch, k0, eps)
    return out


# ---------------------------------------------------------------------------------------------- C6 (compiled program)
def c6_program(n_obs=1000):
    rng = np.random.default_rng(8)
    xs = rng.normal(size=n_obs)
    ys = 1.5 * xs - 0.7 + 0.4 * rng.normal(size=n_obs)
    return [("assume", "a", ["normal", 0.0, 2.0]), ("assume", "b", ["normal", 0.0, 2.0]), ("assume", "s", ["gamma", 2.0, 2.0])] + \
           [("observe", ["normal", ["add", ["mul", "a", float(x)], "b"], "s"], float(y)) for x, y in zip(xs, ys)]


def c6_cpu_baseline(prog, transitions=6):
    """THE REFERENCE'S LITE BACKEND on the same program (oracle/_ref, the py3 rendering): one trace, K transitions of
    (resimulation_mh default one 1) through Trace.primitive_infer -- Lite is single-threaded per trace."""
    ref = os.path.join(ROOT, "oracle", "_ref")
    if not os.path.isdir(os.path.join(ref, "venture", "lite")):
        return None
    import warnings
    warnings.simplefilter("ignore")
    if ref not in sys.path:
        sys.path.insert(0, ref)
    import venture.lite.value as vv
    import venture.value.dicts as vd
    from venture.lite.trace import Trace
    t = Trace(1)
    did = 0

    def sx(e):
        if isinstance(e, str):
            return vd.symbol(e)
        if isinstance(e, (int, float)):
            return vd.number(e)
        return [sx(x) for x in e]
    for d in prog:
        did += 1
        if d[0] == "assume":
            t.eval(did, sx(d[2])); t.bindInGlobalEnv(d[1], did)
        else:
            t.eval(did, sx(d[1])); t.observe(did, vd.number(d[2]))
    t.makeConsistent()
    exp = ["resimulation_mh", vv.VentureSymbol("default"), vv.VentureSymbol("one"), vv.VentureNumber(1)]
    t.primitive_infer(exp)
    t0 = time.perf_counter()
    for _ in range(transitions):
        t.primitive_infer(exp)
    dt = time.perf_counter() - t0
    return {"value": transitions / dt, "unit": "transitions/s", "cores": 1, "kind": "reference",
            "sample": "one Lite trace of the same program, %d transitions of (resimulation_mh default one 1) through "
                      "Trace.primitive_infer (reference Lite backend, py3 rendering), %.1f s" % (transitions, dt)}


def c6(ctx, P=65536, n_obs=1000, K=30, cpu=False):
    """A program nobody wrote a kernel for, through the model compiler (compiler.py -> k_program.cu): Bayesian linear
    regression with unknown noise, 1000 observations, (resimulation_mh default one K)."""
    import venturecxx_b200 as vb
    from venturecxx_b200.compiler import CompiledModel
    prog = c6_program(n_obs)
    t0 = time.perf_counter()
    model = CompiledModel(prog)
    t_compile = time.perf_counter() - t0
    ts = vb.B200TraceSet(model, P, seed=12, device=ctx.local, particle_base=ctx.rank * P)
    ts.use_stream(ctx.stream.cuda_stream)
    ts.incorporate(return_increments=False)
    from venturecxx_b200 import _lib
    run = lambda k: _lib.call("vk_mh", ts._h, 0, _lib.BLOCK_ONE, int(k), None, None)
    run(3)
    ts.synchronize()
    ms = ctx.timed(lambda: run(K), ts.synchronize)
    a = ts.read_site(0)
    ts.close()
    dt = ms * 1e-3
    tr = P * ctx.world * K / dt
    # a site's scaffold: a and b absorb at all n_obs observations, s as well: every transition evaluates them twice (rho, xi)
    evals = tr * 2 * n_obs
    out = {"workload": "compiled program (model compiler, no hand-written kernel): linear regression a, b ~ normal, s ~ gamma, "
                       "%d observations normal(a x_i + b, s), %d particles per GPU, (resimulation_mh default one %d)" % (n_obs, P, K),
           "metric": "particle-site MH transitions/sec", "value": tr, "unit": "transitions/s", "n_gpus": ctx.world, "scaling": "weak",
           "ms_per_step": ms, "obs_loglik_evals_per_sec": evals, "compile_seconds": t_compile, "posterior_mean_a": float(a.mean()),
           "gpu_launches_per_step": 1,
           "roofline": roofline(P * ctx.world * K * (2 * n_obs * 16.0 + 3 * 8.0), dt, ctx.world,
                                "algorithmic bytes: per transition the group table (x_i, y_i: 16 B per observation, twice) + the sites; "
                                "the table is shared by all particles (L2), the path is bound by instruction issue of the interpreter")}
    if cpu and ctx.rank == 0:
        out["cpu_baseline"] = c6_cpu_baseline(prog)
    return out


ALL = {"c1": c1, "c3": c3, "c4": c4, "c5": c5, "c6": c6}


def main():
    import torch
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    quick = "--quick" in sys.argv
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    ctx = Ctx(world, rank, local, stream)
    for w in (args or ["c1", "c3", "c4", "c5", "c6"]):
        kw = {"cpu": "--cpu" in sys.argv}
        if quick and w == "c5":
            kw.update(N=10 ** 6, K=20)
        if quick and w == "c1":
            kw.update(reps=200)
        out = ALL[w](ctx, **kw)
        if rank == 0:
            print(json.dumps({w: out}), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
