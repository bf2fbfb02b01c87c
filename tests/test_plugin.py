"""Plugin surface (venturecxx_b200/plugin.py): __venture_start__ binds the infer
primitives as ripl callbacks with the reference's call_back convention
(python/lib/engine/inference.py:166-177: fn(inferrer, *per-particle-lists))."""
import numpy as np
import pytest

import venturecxx_b200 as vb
from venturecxx_b200 import plugin


class FakeRipl(object):
    def __init__(self):
        self.callbacks = {}

    def bind_callback(self, name, fn):      # ripl.py:941-947
        self.callbacks[name] = fn


class RecordingTraceSet(object):
    def __init__(self):
        self.calls = []

    def primitive_infer(self, exp):
        self.calls.append(exp)
        return np.zeros(3)

    def resample(self, P, mode="multinomial"):
        self.calls.append(("resample", P, mode))

    def likelihood_weight(self):
        self.calls.append("likelihood_weight")


def test_plugin_binds_callbacks_cpu():
    r, ts = FakeRipl(), RecordingTraceSet()
    p = plugin.__venture_start__(r, trace_set=ts)
    assert p.trace_set is ts
    for name in ("b200_resample", "b200_likelihood_weight", "b200_resimulation_mh", "b200_mh", "b200_subsampled_mh",
                 "b200_incorporate", "b200_log_likelihood_at"):
        assert name in r.callbacks
    # call_back passes one list per argument (the value in every particle); stack dicts are unboxed
    out = r.callbacks["b200_resimulation_mh"](None, ["default"] * 3, [{"type": "symbol", "value": "one"}] * 3, [7] * 3)
    assert out == [0.0, 0.0, 0.0]
    assert ts.calls[-1] == ["resimulation_mh", "default", "one", 7]
    r.callbacks["b200_subsampled_mh"](None, "default", "one", 64, 3, 0.01, False, 0, False, 5)
    assert ts.calls[-1] == ["subsampled_mh", "default", "one", 64, 3, 0.01, False, 0, False, 5]
    r.callbacks["b200_resample"](None, [8])
    assert ts.calls[-1] == ("resample", 8, "multinomial")
    with pytest.raises(vb.VentureException):
        plugin.__venture_start__(FakeRipl())


@pytest.mark.gpu
def test_plugin_runs_on_device():
    r = FakeRipl()
    p = plugin.__venture_start__(r, model=vb.NormalNormalModel(0.0, 1.0, 1.0), n_particles=4096, seed=5)
    p.trace_set.observe([2.0])
    r.callbacks["b200_incorporate"](None)
    r.callbacks["b200_likelihood_weight"](None)
    r.callbacks["b200_resample"](None, [4096])
    nodes = r.callbacks["b200_mh"](None, ["default"], ["one"], [20])
    assert len(nodes) == 4096
    x = p.trace_set.read_site(0)
    # posterior N(1, sqrt(1/2))  (test/smc/test_particle_filter.py:61-72)
    assert abs(x.mean() - 1.0) < 0.06 and abs(x.std() - np.sqrt(0.5)) < 0.05
