"""Times one vk_subsampled_mh call (C5 shape) and prints a digest of its decisions.
Usage: python tools/diag_ssp.py N K [P]; knobs come from the environment (VK_SS_HOST_LOOP,
VK_SS_GRID, VK_SS_ROUNDS_PER_BATCH, VK_SS_BATCH_WORK)."""
import hashlib
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import venturecxx_b200 as vb  # noqa: E402
from tools.bench_configs import c5_data  # noqa: E402


def main():
    N, K = int(sys.argv[1]), int(sys.argv[2])
    P = int(sys.argv[3]) if len(sys.argv) > 3 else 64
    D = 16
    X, y = c5_data(N, D)
    ts = vb.B200TraceSet(vb.LogisticRegressionModel(D, N), P, seed=9)
    ts.observe(X, y)
    ts.incorporate(return_increments=False)
    ts.rng_step = 1000
    ts.subsampled_mh("default", "one", 65536, 3, 0.01, False, 0, False, 2)
    ts.synchronize()
    ts.rng_step = 2000
    t0 = time.perf_counter()
    st = ts.subsampled_mh("default", "one", 65536, 3, 0.01, False, 0, False, K)
    ts.synchronize()
    dt = time.perf_counter() - t0
    w = np.stack([ts.read_site(j) for j in range(D)])
    print(json.dumps({"N": N, "K": K, "P": P, "seconds": dt, "transitions_per_s": P * K / dt, "ms_per_transition": 1e3 * dt / K,
                      "accepted": float(st[:, 0].sum()), "mean_sections": float(st[:, 1].mean()),
                      "digest": hashlib.sha1(np.ascontiguousarray(st[:, :2]).tobytes() + w.tobytes()).hexdigest()[:12],
                      "env": {k: v for k, v in os.environ.items() if k.startswith("VK_SS")}}), flush=True)


if __name__ == "__main__":
    main()
