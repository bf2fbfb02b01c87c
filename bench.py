#!/usr/bin/env python
"""bench.py -- headline benchmark of the particle-parallel inference path.

Workload (BASELINE.json configs[1], SURVEY 8(d) D2 "C2"): Bayesian logistic
regression, D = 16 weights, N = 10^4 synthetic observations, 65 536 particles
per GPU, program `(resimulation_mh default one 160)` (10 sweeps) per step.

  python bench.py [--gpus N] [--steps K] [--warmup W]
  python bench.py --impl reference ...     # CPU arm: the reference's Lite backend (oracle/_ref) on all host cores

One JSON line on rank 0.  `value` = particle-site MH transitions/s with the
particle state resident in HBM (device-timed, max over ranks); `e2e` = the same
metric through the public host-buffer API (forget + observe(X, y) H2D +
incorporate + resimulation_mh + D2H of the node counts, per step).
`configs` carries the other BASELINE.json configs at the same GPU count
(tools/bench_configs.py): c1 normal-normal, c3 HMM filter with the particles
sharded over the GPUs (peer-memory resample), c4 LDA, c5 subsampled_mh with the
observations sharded over the GPUs (peer mailbox) -- each with its own roofline,
identity flag against the 1-GPU result and, at one GPU, a cpu_baseline.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "particle-site MH transitions/sec"
UNIT = "transitions/s"


def make_data(N, D):
    # X ~ N(0,1) (seed 0), w* ~ N(0,1) (seed 1), y ~ Bernoulli(logistic(x.w*)) (seed 2)
    X = np.random.default_rng(0).normal(size=(N, D))
    w_true = np.random.default_rng(1).normal(size=D)
    y = (np.random.default_rng(2).random(N) < 1.0 / (1.0 + np.exp(-X @ w_true))).astype(np.uint8)
    return np.ascontiguousarray(X), y


def load_traffic(P, N):
    """Measured DRAM bytes per launch of the dominant kernel (ncu --set full capture committed under
    profiles/), scaled to this run's particle count; None if the capture does not match the workload."""
    try:
        with open(os.path.join(ROOT, "profiles", "r2_blr_mh2_traffic.json")) as f:
            t = json.load(f)
        if int(t["observations"]) != int(N):
            return None, None
        return (t["dram_bytes_read"] + t["dram_bytes_write"]) * (float(P) / t["particles"]), t
    except Exception:
        return None, None


def fp64_frac(tinfo, P, N, T, ms_per_step, clocks):
    """fp64 warp instructions per second of the timed launch over the sustained DFMA rate of the GPU"""
    try:
        evals = float(P) * T * (-(-N // 1024) * 1024) / 32.0                 # warp-evaluations (rows padded to the kernel's tile)
        share = tinfo["hot_loop_fp64_instr_per_eval"] * evals                 # hot loop only: a lower bound of the launch's fp64 count
        mhz = (clocks or {}).get("sm_mhz") or 1965.0
        peak = tinfo["measured_dfma_peak_warp_instr_per_clk_per_sm"] * 148 * mhz * 1e6
        return share / (ms_per_step * 1e-3) / peak
    except Exception:
        return None


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(object):
    """nvidia-smi clocks/throttle reasons DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def cpu_baseline(X, y, D, N, T, seconds_target=12.0):
    """Oracle (CPU restatement of the Lite algorithm) on all host cores, bounded sample."""
    import oracle
    oracle.build()
    L = oracle.lib()
    cores = os.cpu_count() or 1

    def run(Ps, Ts):
        w = np.ascontiguousarray(np.random.default_rng(3).normal(size=(D, Ps)))
        t0 = time.perf_counter()
        L.orc_blr_mh(oracle._ptr(w), oracle._ptr(X), oracle._ptr(y), oracle.I32(D), oracle.I64(N), oracle.I64(Ps),
                     oracle.I64(0), oracle.F64(0.0), oracle.F64(1.0), oracle.U64(1), oracle.U32(0), oracle.I32(Ts),
                     oracle.I32(-1), None, None)
        return time.perf_counter() - t0

    Ts = min(T, 16)
    dt = run(cores, Ts)                      # calibration: one particle per core
    per = dt / Ts                            # seconds per transition per core-slot
    Ps = int(max(cores, min(4096, (seconds_target / max(per * Ts, 1e-9)) * cores)))
    Ps -= Ps % cores
    dt = run(Ps, Ts)
    return {"value": Ps * Ts / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": "%d particles x %d transitions of the same N=%d, D=%d workload, OpenMP over particles, %.1f s"
                      % (Ps, Ts, N, D, dt)}, dt


# ---- reference arm: the reference's own Lite backend (oracle/_ref) on the host cores -------------------
_LITE = {}


def _lite_worker_init(X, y, D, N, seed_base):
    """One Lite trace of the C2 model per worker process (Lite is single-threaded per trace; the reference
    parallelises over traces with worker processes, python/lib/multiprocess.py)."""
    import warnings
    warnings.simplefilter("ignore")
    sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
    import venture.lite.value as vv
    import venture.value.dicts as vd
    from venture.lite.trace import Trace
    t = Trace(seed_base + os.getpid() % 100003)
    S, n = vd.symbol, vd.number
    for j in range(D):
        t.eval(j + 1, [S("normal"), n(0.0), n(1.0)])
        t.bindInGlobalEnv("w_%d" % j, j + 1)
    t.eval(D + 1, [S("array")] + [S("w_%d" % j) for j in range(D)])
    t.bindInGlobalEnv("w", D + 1)
    for i in range(N):
        did = D + 2 + i
        t.eval(did, [S("log_odds_flip"), [S("vector_dot"), S("w"), [S("array")] + [n(float(v)) for v in X[i]]]])
        t.observe(did, vd.boolean(bool(y[i])))
    t.makeConsistent()
    _LITE["trace"] = t
    _LITE["exp"] = lambda k: ["resimulation_mh", vv.VentureSymbol("default"), vv.VentureSymbol("one"), vv.VentureNumber(k)]


def _lite_worker_step(k):
    t0 = time.perf_counter()
    _LITE["trace"].primitive_infer(_LITE["exp"](k))     # infer/dispatch.py:130-136, the stock code path
    return time.perf_counter() - t0


def _lite_worker_ready(_):
    return "trace" in _LITE


def run_reference_lite(args, X, y, D, N, T):
    """`(resimulation_mh default one k)` on one real Lite trace per host core, N and D as in the GPU arm.
    A step is k = 1 transition per trace (a Lite transition of this model walks 2N + 3 nodes: ~3 s)."""
    import multiprocessing as mp
    cores = os.cpu_count() or 1
    try:                                    # a Lite trace of this model holds ~0.36 GB: stay inside the box's memory
        with open("/proc/meminfo") as f:
            avail_kb = [int(ln.split()[1]) for ln in f if ln.startswith("MemAvailable")][0]
        cores = max(1, min(cores, int(avail_kb / 1024 / 1024 / 0.75)))
    except Exception:
        cores = min(cores, 16)
    cores = min(cores, 64)
    ctx = mp.get_context("fork")
    t_build = time.perf_counter()
    pool = ctx.Pool(cores, initializer=_lite_worker_init, initargs=(X, y, D, N, 1000))
    assert all(pool.map(_lite_worker_ready, range(cores), chunksize=1))
    t_build = time.perf_counter() - t_build
    k = 1
    times = []
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        pool.map(_lite_worker_step, [k] * cores, chunksize=1)
        dt = time.perf_counter() - t0
        if i >= args.warmup:
            times.append(dt)
    pool.close(); pool.join()
    tot = sum(times)
    value = cores * k * len(times) / tot
    cb = {"value": value, "unit": UNIT, "cores": cores, "kind": "reference",
          "sample": "%d Lite traces (one per core, py3 rendering of backend/lite in oracle/_ref) x %d transition(s) per "
                    "step of the same N=%d, D=%d model; traces built in %.1f s outside the timed region"
                    % (cores, k, N, D, t_build)}
    return value, 1e3 * tot / max(len(times), 1), cb, \
        "the reference's Lite backend itself (backend/lite, rendered importable under Python 3 by " \
        "oracle/lite_py3/build_lite_py3.py): scaffold construction + detach + regen over 2N+3 nodes per transition"


def run_reference(args, X, y, D, N, T):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if os.path.isdir(os.path.join(ROOT, "oracle", "_ref", "venture", "lite")):
        value, ms, cb, note = run_reference_lite(args, X, y, D, N, T)
    else:
        vals = []
        for i in range(args.warmup + args.steps):
            cb, dt = cpu_baseline(X, y, D, N, T, seconds_target=4.0)
            if i >= args.warmup:
                vals.append((cb, dt))
        value = float(np.mean([cb["value"] for cb, _ in vals]))
        ms = 1e3 * sum(dt for _, dt in vals) / max(len(vals), 1)
        cb = dict(vals[-1][0]); cb["value"] = value
        note = ("oracle/_ref (the py3 rendering of the reference's Lite backend) is not present on this box; this arm "
                "times the oracle port (C restatement of the Lite algorithm, OpenMP over particles)")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "C2 Bayesian logistic regression, single-site resimulation_mh (default one)",
                   "observations": N, "D": D,
                   "timed": "one trace per host core (%d traces) x 1 transition per step; the GPU arm's 65 536 particles x 160 "
                            "transitions per step would take this arm weeks, the rate per transition is what is compared"
                            % cb["cores"],
                   "particles": cb["cores"], "transitions_per_step": 1},
        "cpu_baseline": cb,
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "obs_loglik_evals_per_sec": value * N,
        "note": note,
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--particles", type=int, default=65536, help="particles per GPU")
    ap.add_argument("--obs", type=int, default=10000)
    ap.add_argument("--dim", type=int, default=16)
    ap.add_argument("--transitions", type=int, default=160)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the secondary configs (c1, c3, c4, c5, c6)")
    ap.add_argument("--configs", default="c1,c3,c4,c5,c6")
    ap.add_argument("--configs-timeout", type=int, default=420, help="seconds before the secondary configs are abandoned")
    args = ap.parse_args()
    D, N, T, P = args.dim, args.obs, args.transitions, args.particles
    X, y = make_data(N, D)

    if args.impl == "reference":
        run_reference(args, X, y, D, N, T)
        return

    import torch
    import torch.distributed as dist
    import venturecxx_b200 as vb

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    W = max(args.warmup, 3)

    ts = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), P, seed=1234, device=local, particle_base=rank * P)
    stream = torch.cuda.Stream()          # a real (non-default) stream shared by torch events and the library
    torch.cuda.set_stream(stream)
    ts.use_stream(stream.cuda_stream)
    ts.observe(X, y)
    ts.incorporate()
    from venturecxx_b200 import _lib

    def step_resident():
        # state resident in HBM, no host outputs: exactly one kernel launch
        _lib.call("vk_mh", ts._h, 0, _lib.BLOCK_ONE, T, None, None)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # L2 note: the per-particle logit rows total P*N*8 B = 5.2 GB >> 126 MB L2; Xs (1.3 MB) is meant to be L2 resident.
    for _ in range(W):
        step_resident()
    acc0 = np.empty(P, dtype=np.int64)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record(stream)
    for _ in range(args.steps):
        step_resident()
    ev1.record(stream)
    barrier()
    ms = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if rank == 0 else None
    # accept rate of one extra (untimed) step, for the algorithmic-bytes formula
    _lib.call("vk_mh", ts._h, 0, _lib.BLOCK_ONE, T, None, _lib.ptr(acc0))
    a_rate = float(acc0.mean()) / T

    # ---- end to end through the public API with host buffers ------------
    Xp = torch.from_numpy(X).pin_memory().numpy()
    yp = torch.from_numpy(y).pin_memory().numpy()

    nodes_pinned = torch.empty(P, dtype=torch.float64).pin_memory().numpy()
    phases = {"forget": 0.0, "observe_h2d": 0.0, "incorporate": 0.0, "resimulation_mh_d2h": 0.0}

    def step_e2e():
        # the program a user runs per batch of data: forget the old observations, observe the new ones (HOST
        # buffers), incorporate, infer, read the result back.  Host time per phase is recorded (the calls with
        # host buffers drain the stream, so the last phase carries the kernels' time).
        t = time.perf_counter()
        ts.forget_observations(return_weights=False)
        t1 = time.perf_counter(); phases["forget"] += t1 - t
        ts.observe(Xp, yp)                       # H2D: N*D*8 + N bytes (pinned)
        t2 = time.perf_counter(); phases["observe_h2d"] += t2 - t1
        ts.incorporate(return_increments=False)  # stream-ordered
        t3 = time.perf_counter(); phases["incorporate"] += t3 - t2
        nodes = ts.resimulation_mh("default", "one", T, out=nodes_pinned)   # D2H: P*8 into pinned memory
        phases["resimulation_mh_d2h"] += time.perf_counter() - t3
        return nodes

    step_e2e()
    for k in phases:
        phases[k] = 0.0
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record(stream)
    for _ in range(args.steps):
        step_e2e()
    e1.record(stream)
    barrier()
    ms_e2e = max(e0.elapsed_time(e1), 1e3 * (time.perf_counter() - t0))

    # ---- fp32 throughput mode (vk_set_precision): reported beside the headline, never as it -----
    ts.precision = "fp32"
    for _ in range(W):
        step_resident()
    barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record(stream)
    for _ in range(args.steps):
        step_resident()
    f1.record(stream)
    barrier()
    ms_f32 = f0.elapsed_time(f1)
    ts.precision = "fp64"

    t_ms = torch.tensor([ms, ms_e2e, ms_f32], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
    ms, ms_e2e, ms_f32 = [float(v) for v in t_ms.tolist()]
    ts.close()

    out = None
    if rank == 0:
        total_tr = float(P) * world * T * args.steps
        value = total_tr / (ms * 1e-3)
        e2e_value = total_tr / (ms_e2e * 1e-3)
        peak, peak_src = load_peaks()
        # N*s*(1+a) + 2s bytes per particle-transition (s = 8)
        bytes_per_tr = N * 8.0 * (1.0 + a_rate) + 16.0
        launch_s = ms * 1e-3 / args.steps
        achieved = bytes_per_tr * P * T / launch_s / 1e9
        traffic, tinfo = load_traffic(P, N)
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": W,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": "C2 Bayesian logistic regression, single-site resimulation_mh (default one)",
                       "particles_per_gpu": P, "observations": N, "D": D, "transitions_per_step": T,
                       "l2": "inputs larger than L2 (logit rows %.1f GB per GPU)" % (P * N * 8 / 1e9)},
            "obs_loglik_evals_per_sec": value * N,
            "accept_rate": a_rate,
            "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": ms_e2e / args.steps,
                    "h2d_bytes_per_step": int(X.nbytes + y.nbytes), "d2h_bytes_per_step": int(P * 8),
                    "host_ms_per_phase": dict((k, 1e3 * v / args.steps) for k, v in phases.items())},
            "gpu_launches": args.steps,
            "fp32_mode": {"value": total_tr / (ms_f32 * 1e-3), "unit": UNIT, "ms_per_step": ms_f32 / args.steps,
                          "note": "same workload and steps with vk_set_precision(VK_FP32): float logit row and MUFU "
                                  "log-logistic inside the MH loop, fp64 proposals/accept test/state; log densities "
                                  "within 1e-5 relative of fp64 (tests/test_gpu_fp32_mode.py). Not the headline: "
                                  "`value`, `e2e` and `roofline` are the fp64 path."},
            "clocks": clocks,
            "configs": {},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                         "kernel": "blr_mh2_kernel<2 CTAs/SM> (256 threads, two proposals per pass over the shared-memory logit row, 8 chains per thread)",
                         "algorithmic_bytes_per_launch": bytes_per_tr * P * T,
                         "pipes_from_ncu": None if tinfo is None else {
                             "fp64_pipe_pct": tinfo["fp64_pipe_pct"], "issue_active_pct": tinfo["issue_active_pct"],
                             "lsu_writeback_pct": tinfo["lsu_writeback_pct"]},
                         # the bound that binds (live): fp64 warp instructions of the launch / time, against the DFMA rate
                         # one SM SUSTAINS (tools/ubench/fp64_pipe.cu on this pool: 1.75 warp-instr/clk/SM of the nominal 2)
                         "fp64_pipe_frac": None if tinfo is None else fp64_frac(tinfo, P, N, T, ms / args.steps, clocks),
                         "issue_slot_model": None if tinfo is None else {
                             "slots_per_eval": 2 * tinfo["hot_loop_fp64_instr_per_eval"] + (tinfo["hot_loop_instr_per_eval"] - tinfo["hot_loop_fp64_instr_per_eval"]),
                             "note": "a DFMA occupies two issue slots of its SM sub-partition (micro-benchmark: DFMA alone 1.75/clk/SM, "
                                     "with one IMAD per DFMA 1.16, with IMAD + LOP3 0.88); per evaluation the hot loop issues 12 fp64 + "
                                     "10.25 other instructions = 34.25 slots, i.e. 39 cycles per warp-evaluation per sub-partition "
                                     "at the sustained 87 %; measured 46.6 in the hot loop (profiles/r2_blr_mh2_ncu.md)"},
                         "note": "algorithmic bytes per SURVEY 8(d) D4 = N*8*(1+a)+16 per particle-transition; "
                                 "the kernel keeps the logit row in shared memory for all transitions of a launch, so "
                                 "real DRAM traffic (`traffic`, from the committed ncu capture) is 2*8N B per particle per LAUNCH, "
                                 "~1/200 of the algorithmic figure; the binding resource is issue slots, of which every "
                                 "fp64 instruction takes two (profiles/r2_blr_mh2_ncu.md, profiles/r2_fp64_pipe_microbench.txt)"},
        }
        if not args.no_cpu_baseline and world == 1:
            cb, _ = cpu_baseline(X, y, D, N, T)
            out["cpu_baseline"] = cb

    # ---- the other BASELINE.json configs at this GPU count (tools/bench_configs.py) ---------------
    # The headline line must survive whatever happens here: a watchdog on EVERY rank prints it (rank 0) and
    # leaves after --configs-timeout seconds (a rank stuck in a collective cannot be recovered otherwise).
    printed = threading.Lock()

    def emit_line():
        if rank == 0 and printed.acquire(False):
            print(json.dumps(out), flush=True)

    def watchdog():
        if out is not None:
            out["configs"]["error"] = "secondary configs did not finish within %d s; records so far kept" % args.configs_timeout
        emit_line()
        os._exit(0)

    if not args.no_configs:
        timer = threading.Timer(args.configs_timeout, watchdog)
        timer.daemon = True
        timer.start()
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import bench_configs as bc
        ctx = bc.Ctx(world, rank, local, stream)
        want_cpu = (not args.no_cpu_baseline) and world == 1
        for name in [c for c in args.configs.split(",") if c]:
            t_sub = time.perf_counter()
            try:
                rec = bc.ALL[name](ctx, cpu=want_cpu)
                rec["bench_seconds"] = time.perf_counter() - t_sub
            except Exception as e:           # a secondary record never takes the headline line down with it
                import traceback
                rec = {"error": "%s: %s" % (type(e).__name__, e), "traceback": traceback.format_exc()[-1500:]}
                sys.stderr.write("bench config %s failed on rank %d: %s\n" % (name, rank, rec["traceback"]))
            if out is not None:
                out["configs"][name] = rec
        timer.cancel()
    emit_line()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
