#!/usr/bin/env python
"""Secondary measurements: the five BASELINE.json configs on ONE GPU at the
sizes of SURVEY 8(d) D2, one JSON line each (bench.py stays the headline, C2).

  python tools/bench_configs.py [c1 c1big c2 c2f32 c2lw c3 c4 c4gibbs c3gibbs c5] [--quick]

Timed through the public host API (B200TraceSet), wall clock around a
synchronised region; algorithmic bytes per SURVEY 8(d) D4.
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import venturecxx_b200 as vb  # noqa: E402

PEAK = 6538.9
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass


def emit(name, metric, value, unit, seconds, bytes_moved=None, **extra):
    out = {"config": name, "metric": metric, "value": value, "unit": unit, "seconds": seconds}
    if bytes_moved is not None:
        gbs = bytes_moved / seconds / 1e9
        out["roofline"] = {"bound": "hbm", "achieved": gbs, "peak": PEAK, "unit": "GB/s", "frac": gbs / PEAK}
    out.update(extra)
    print(json.dumps(out), flush=True)


def c1(P=1000, reps=1000, name="c1"):
    """(do (resample P) (likelihood_weight) (resample P)) x reps, normal-normal (benchmarks/Makefile:6-9)"""
    ts = vb.B200TraceSet(vb.NormalNormalModel(0.0, 1.0, 1.0), P, seed=1)
    ts.observe([2.0])
    ts.incorporate()
    for _ in range(5):
        ts.resample(P, mode="systematic"); ts.likelihood_weight(); ts.resample(P, mode="systematic")
    ts.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        ts.resample(P, mode="systematic"); ts.likelihood_weight(); ts.resample(P, mode="systematic")
    ts.synchronize()
    dt = time.perf_counter() - t0
    # D4: likelihood_weight 2s, resample 2s + 8 + 2*state(8+8) per particle, twice
    b = reps * P * (16 + 2 * (16 + 8 + 32))
    emit(name, "particle program iterations/sec (resample, likelihood_weight, resample)", reps * P / dt,
         "particle-iterations/s", dt, b, particles=P, reps=reps, us_per_rep=1e6 * dt / reps)


def c3(P=1 << 20, quick=False):
    H = json.load(open(os.path.join(ROOT, "tests", "golden", "hmm.json")))
    ts = vb.B200TraceSet(vb.HMMModel(H["transition_matrix"], H["emission_matrix"], 1, 20), P, seed=5)

    def run():
        tv = []
        for t, d in enumerate(H["data"]):
            ts.observe(d); ts.incorporate(return_increments=False); ts.resample(P, mode="systematic")
        return tv
    run()
    ts.close()
    ts = vb.B200TraceSet(vb.HMMModel(H["transition_matrix"], H["emission_matrix"], 1, 20), P, seed=6)
    ts.synchronize()
    t0 = time.perf_counter()
    run()
    ts.synchronize()
    dt = time.perf_counter() - t0
    T = len(H["data"])
    m = np.bincount(ts.read_site(T - 1).astype(int), minlength=6)[1:6] / P
    tvd = 0.5 * np.abs(m - np.array(H["filtered_marginals"][T - 1])).sum()
    emit("c3", "HMM particle-filter particle-steps/sec (observe, incorporate, resample)", P * T / dt,
         "particle-steps/s", dt, P * T * 208.0, particles=P, steps=T, ms_per_step=1e3 * dt / T, tvd_last=tvd)


def lda_corpus(K, V, n_docs, n_per, seed=3):
    rng = np.random.default_rng(seed)
    theta = rng.dirichlet(np.full(K, 0.5), n_docs)
    phi = rng.dirichlet(np.full(V, 0.5), K)
    doc, word = [], []
    for d in range(n_docs):
        for k in rng.choice(K, n_per, p=theta[d]):
            doc.append(d); word.append(rng.choice(V, p=phi[k]))
    return np.array(doc, dtype=np.int32), np.array(word, dtype=np.int32)


def c4(P=4096, sweeps=4, quick=False):
    K, V, n_docs, n_per = 16, 1024, 64, 64
    doc, word = lda_corpus(K, V, n_docs, n_per)
    n_tok = len(doc)
    ts = vb.B200TraceSet(vb.LDAModel(K, V, n_docs, n_tok), P, seed=7)
    ts.observe(doc, word)
    ts.incorporate()
    n_sites = n_tok + 2
    ts.primitive_infer(["resimulation_mh", "default", "one", n_sites])
    ts.synchronize()
    t0 = time.perf_counter()
    for _ in range(sweeps):
        ts.primitive_infer(["resimulation_mh", "default", "one", n_sites])
    ts.synchronize()
    dt = time.perf_counter() - t0
    emit("c4", "LDA particle-site MH transitions/sec (resimulation_mh default one)", P * n_sites * sweeps / dt,
         "transitions/s", dt, P * n_sites * sweeps * 100.0, particles=P, sites=n_sites, sweeps=sweeps)


def c4gibbs(P=4096, sweeps=4, quick=False):
    """(gibbs z each 1) sweeps: enumerative Gibbs over the 4096 z sites."""
    K, V, n_docs, n_per = 16, 1024, 64, 64
    doc, word = lda_corpus(K, V, n_docs, n_per)
    n_tok = len(doc)
    ts = vb.B200TraceSet(vb.LDAModel(K, V, n_docs, n_tok), P, seed=7)
    ts.observe(doc, word)
    ts.incorporate()
    ts.primitive_infer(["gibbs", "z", "each", 1])
    ts.synchronize()
    t0 = time.perf_counter()
    ts.primitive_infer(["gibbs", "z", "each", sweeps])
    ts.synchronize()
    dt = time.perf_counter() - t0
    # D4-style algorithmic bytes: z rw 2 + doc row K*4 + K topic-word cells K*4 + K totals K*4 + 2 cell updates 16 + ids 4
    emit("c4gibbs", "LDA enumerative-Gibbs z-site transitions/sec (gibbs z each)", P * n_tok * sweeps / dt,
         "transitions/s", dt, P * n_tok * sweeps * (12.0 * K + 22.0), particles=P, sites=n_tok, sweeps=sweeps)


def c3gibbs(P=1 << 20, sweeps=4, quick=False):
    """(gibbs default each 1) sweeps over state(2..20) of hmm.vnt after the 20 observations."""
    H = json.load(open(os.path.join(ROOT, "tests", "golden", "hmm.json")))
    ts = vb.B200TraceSet(vb.HMMModel(H["transition_matrix"], H["emission_matrix"], 1, 20), P, seed=5)
    for d in H["data"]:
        ts.observe(d); ts.incorporate(return_increments=False)
    ts.primitive_infer(["gibbs", "default", "each", 1])
    ts.synchronize()
    t0 = time.perf_counter()
    ts.primitive_infer(["gibbs", "default", "each", sweeps])
    ts.synchronize()
    dt = time.perf_counter() - t0
    T = len(H["data"])
    # per site: 3 state bytes read + 1 written; the history row (32 B) is read and written once per sweep
    emit("c3gibbs", "HMM enumerative-Gibbs state-site transitions/sec (gibbs default each)", P * (T - 1) * sweeps / dt,
         "transitions/s", dt, P * sweeps * 64.0, particles=P, sites=T - 1, sweeps=sweeps)


def c2(precision="fp64", P=65536, N=10000, D=16, K=160, reps=3, quick=False):
    """C2 through the host API at either precision (bench.py is the headline and always fp64)."""
    if quick:
        P, reps = 8192, 2
    X = np.random.default_rng(0).normal(size=(N, D))
    w_true = np.random.default_rng(1).normal(size=D)
    y = (np.random.default_rng(2).random(N) < 1.0 / (1.0 + np.exp(-X @ w_true))).astype(np.uint8)
    ts = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), P, seed=11, precision=precision)
    ts.observe(X, y)
    ts.incorporate(return_increments=False)
    ts.resimulation_mh("default", "one", K)
    ts.synchronize()
    t0 = time.perf_counter()
    acc = 0
    for _ in range(reps):
        _, a = ts.resimulation_mh("default", "one", K, return_accepts=True)
        acc += int(a.sum())
    ts.synchronize()
    dt = time.perf_counter() - t0
    tr = float(P) * K * reps
    emit("c2" + ("f32" if precision == "fp32" else ""), "particle-site MH transitions/sec", tr / dt, "transitions/s", dt,
         tr * (N * 8 * (1 + acc / tr) + 16), precision=precision, particles=P, observations=N, D=D,
         obs_evals_per_sec=tr * N / dt, accept_rate=acc / tr)


def c2lw(P=65536, N=10000, D=16, reps=5, quick=False):
    """C2 likelihood_weight / incorporate: observation log-likelihood evaluations per second."""
    if quick:
        P = 8192
    X = np.random.default_rng(0).normal(size=(N, D))
    y = (np.random.default_rng(2).random(N) < 0.5).astype(np.uint8)
    ts = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), P, seed=11)
    ts.observe(X, y)
    ts.incorporate(return_increments=False)
    ts.likelihood_weight()
    ts.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        ts.likelihood_weight()
    ts.synchronize()
    dt = time.perf_counter() - t0
    ev = float(P) * N * reps
    emit("c2lw", "observation log-lik evals/sec (likelihood_weight)", ev / dt, "evals/s", dt, ev * 8.0 + reps * P * (D + 2) * 8.0,
         particles=P, observations=N, D=D, ms_per_call=1e3 * dt / reps)


def c5(N=10 ** 7, P=64, K=20, quick=False):
    D = 16
    if quick:
        N = 10 ** 6
    X = np.random.default_rng(4).normal(size=(N, D))
    w_true = np.random.default_rng(5).normal(size=D)
    y = (np.random.default_rng(6).random(N) < 1.0 / (1.0 + np.exp(-X @ w_true))).astype(np.uint8)
    ts = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), P, seed=9)
    t0 = time.perf_counter()
    ts.observe(X, y)
    ts.incorporate()
    t_obs = time.perf_counter() - t0
    ts.primitive_infer(["subsampled_mh", "default", "one", 65536, 3, 0.01, False, 0, False, 2])
    ts.synchronize()
    t0 = time.perf_counter()
    stats = ts.primitive_infer(["subsampled_mh", "default", "one", 65536, 3, 0.01, False, 0, False, K])
    ts.synchronize()
    dt = time.perf_counter() - t0
    visited = float(stats[:, 1].sum()) * K          # mean sections per transition x transitions, summed over particles
    emit("c5", "subsampled_mh observation log-lik evals/sec", 2 * visited / dt, "evals/s", dt,
         visited * (D * 8 + 1) / 32.0, particles=P, observations=N, transitions=K,
         mean_sections_per_transition=float(stats[:, 1].mean()), accepted=float(stats[:, 0].mean()),
         observe_incorporate_seconds=t_obs, transitions_per_sec=P * K / dt)


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    quick = "--quick" in sys.argv
    which = args or ["c1", "c1big", "c3", "c4", "c4gibbs", "c3gibbs", "c5"]
    for w in which:
        if w == "c1":
            c1(1000, 200 if quick else 1000)
        elif w == "c1big":
            c1(1 << 24, 5 if quick else 20, name="c1big")
        elif w == "c3":
            c3(1 << 20, quick)
        elif w == "c4":
            c4(4096, 1 if quick else 4, quick)
        elif w == "c4gibbs":
            c4gibbs(4096, 1 if quick else 4, quick)
        elif w == "c3gibbs":
            c3gibbs(1 << 20, 1 if quick else 4, quick)
        elif w == "c5":
            c5(quick=quick)
        elif w == "c2lw":
            c2lw(quick=quick)
        elif w == "c2":
            c2("fp64", quick=quick)
        elif w == "c2f32":
            c2("fp32", quick=quick)


if __name__ == "__main__":
    main()
