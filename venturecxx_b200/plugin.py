"""Venture plugin for the B200 particle path.

Loaded the way the reference loads plugins: `ripl.load_plugin(path, *args)`
imports the file and calls `__venture_start__(ripl, *args)`
(python/lib/ripl/ripl.py:1093-1096).  Inside, a plugin registers callbacks
with `ripl.bind_callback(name, fn)` (ripl.py:941-947); the engine invokes
them from an inference program as `(call_back name args...)` with the
signature `fn(inferrer, *args)` where every arg is the list of that
expression's value in each particle (python/lib/engine/inference.py:166-177).

The callbacks forward to ONE device-resident B200TraceSet with the argument
meaning of the reference's inference SPs:
  resample(particles)                         inference_sps.py:536-551
  likelihood_weight()                         inference_sps.py:602-606
  resimulation_mh / mh(scope, block, transitions)        :234-241
  gibbs(scope, block, transitions)                       :253-267
  subsampled_mh(scope, block, Nbatch, k0, epsilon, useDeltaKernels,
                deltaKernelArgs, updateValues, transitions)   :678-694
  incorporate(), log_likelihood_at / log_joint_at(scope, block)
"""
from .exception import VentureException
from .trace_set import B200TraceSet

PREFIX = "b200_"


def _scalar(v):
    """call_back hands every argument over as a per-particle list; inference
    arguments are deterministic, so the first entry is the value."""
    if isinstance(v, (list, tuple)):
        if not v:
            raise VentureException("invalid_argument", "empty argument list")
        v = v[0]
    if isinstance(v, dict) and "value" in v:
        v = v["value"]
    return v


class B200Plugin(object):
    """Holds the particle set and exposes the infer primitives as callbacks."""

    def __init__(self, model, n_particles, seed=1, device=0, trace_set=None):
        self.trace_set = trace_set if trace_set is not None else B200TraceSet(model, n_particles, seed, device)

    # every public method below is bound as `b200_<name>`
    def resample(self, _inferrer, particles, mode="multinomial"):
        self.trace_set.resample(int(_scalar(particles)), mode=_scalar(mode))

    def resample_systematic(self, _inferrer, particles):
        self.trace_set.resample(int(_scalar(particles)), mode="systematic")

    def likelihood_weight(self, _inferrer):
        self.trace_set.likelihood_weight()

    def incorporate(self, _inferrer):
        return self.trace_set.incorporate().tolist()

    def resimulation_mh(self, _inferrer, scope, block, transitions):
        return self.trace_set.primitive_infer(
            ["resimulation_mh", _scalar(scope), _scalar(block), int(_scalar(transitions))]).tolist()

    mh = resimulation_mh

    def gibbs(self, _inferrer, scope, block, transitions):
        return self.trace_set.primitive_infer(
            ["gibbs", _scalar(scope), _scalar(block), int(_scalar(transitions))]).tolist()

    def subsampled_mh(self, _inferrer, scope, block, Nbatch, k0, epsilon, useDeltaKernels, deltaKernelArgs,
                      updateValues, transitions):
        args = [_scalar(a) for a in (scope, block, Nbatch, k0, epsilon, useDeltaKernels, deltaKernelArgs,
                                     updateValues, transitions)]
        return self.trace_set.primitive_infer(["subsampled_mh"] + args).tolist()

    def log_likelihood_at(self, _inferrer, scope, block):
        return self.trace_set.log_likelihood_at(_scalar(scope), _scalar(block)).tolist()

    def log_joint_at(self, _inferrer, scope, block):
        return self.trace_set.log_joint_at(_scalar(scope), _scalar(block)).tolist()

    def particle_log_weights(self, _inferrer):
        return self.trace_set.log_weights.tolist()

    def callbacks(self):
        names = ["resample", "resample_systematic", "likelihood_weight", "incorporate", "resimulation_mh", "mh",
                 "gibbs", "subsampled_mh", "log_likelihood_at", "log_joint_at", "particle_log_weights"]
        return dict((PREFIX + n, getattr(self, n)) for n in names)


def __venture_start__(ripl, model=None, n_particles=1, seed=1, device=0, trace_set=None):
    """Entry point of ripl.load_plugin.  Returns the plugin object so the caller
    can observe data through `plugin.trace_set.observe(...)`."""
    if model is None and trace_set is None:
        raise VentureException("invalid_argument", "the b200 plugin needs a compiled model (venturecxx_b200.models)")
    plugin = B200Plugin(model, n_particles, seed, device, trace_set)
    for name, fn in plugin.callbacks().items():
        ripl.bind_callback(name, fn)
    return plugin
