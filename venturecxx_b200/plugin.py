"""Venture plugin for the B200 particle path.

Loaded the way the reference loads plugins: `ripl.load_plugin(path, *args)`
imports the file and calls `__venture_start__(ripl, *args)`
(python/lib/ripl/ripl.py:1093-1096).  Inside, the plugin

  * binds FOREIGN INFERENCE SPs with `ripl.bind_foreign_inference_sp(name, sp)`
    (ripl.py:938-939 -> engine/engine.py:56-59) whenever the host Venture's
    `venture.lite.inference_sps` is importable: `b200_resample`,
    `b200_likelihood_weight`, `b200_incorporate`, `b200_resimulation_mh` / `b200_mh`,
    `b200_gibbs`, `b200_slice`, `b200_pgibbs`, `b200_subsampled_mh`, `b200_log_likelihood_at`, `b200_log_joint_at`
    with the SP types of backend/lite/inference_sps.py:234-241, 253-267, 536-551, 602-606,
    678-694 (built with that module's own `sequenced_sp` / `transition_oper_type` /
    `infer_action_maker_type`), so `(b200_resimulation_mh default one 160)` is an
    inference action like the builtin; with `override_builtins=True` the builtin NAMES
    (`resample`, `likelihood_weight`, `resimulation_mh`, `mh`, `gibbs`, `subsampled_mh`, ...)
    are rebound too and user programs run unchanged on the batch;
  * registers callbacks with `ripl.bind_callback(name, fn)` (ripl.py:941-947); the
    engine invokes them from an inference program as `(call_back name args...)` with the
    signature `fn(inferrer, *args)` where every arg is the list of that expression's
    value in each particle (python/lib/engine/inference.py:166-177).

(The other route -- the reference's Engine over the B200 BACKEND, where the builtin
inference SPs reach the batch through TraceSet -- is backend.py.)

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


def _python_value(v):
    """operand of an inference SP (a VentureValue of the host) -> python value"""
    if hasattr(v, "asStackDict"):
        v = v.asStackDict()
    v = _scalar(v)
    if isinstance(v, float) and v == int(v):
        return int(v)
    return v


def inference_sps(plugin):
    """name -> SP for ripl.bind_foreign_inference_sp, or {} when the host Venture is not importable.
    The SP types are the reference's own (backend/lite/inference_sps.py)."""
    try:
        import venture.lite.inference_sps as inf
        import venture.lite.types as t
    except ImportError:
        return {}
    ts = plugin.trace_set

    def transition(op):
        def run(*operands):
            return list(ts.primitive_infer([op] + [_python_value(o) for o in operands]))
        return run

    def subsampled(*operands):
        stats = ts.primitive_infer(["subsampled_mh"] + [_python_value(o) for o in operands])
        nodes = getattr(ts, "last_mean_nodes", None)          # what the reference returns: nodes touched per transition
        if nodes is not None:
            return [float(v) for v in nodes]
        return [float(r[0]) for r in stats] if getattr(stats, "ndim", 1) == 2 else list(stats)

    def resample(particles):
        ts.resample(int(_python_value(particles)), mode="multinomial")

    def likelihood_weight():
        ts.likelihood_weight()

    def incorporate():
        return list(ts.incorporate())

    def ll_at(scope, block):
        return list(ts.log_likelihood_at(_python_value(scope), _python_value(block)))

    def lj_at(scope, block):
        return list(ts.log_joint_at(_python_value(scope), _python_value(block)))

    subsampled_extra = [t.IntegerType("Nbatch : int"), t.IntegerType("k0 : int"), t.NumberType("epsilon : number"),
                        t.BoolType("useDeltaKernels : bool"), t.NumberType("deltaKernelArgs : number"),
                        t.BoolType("updateValues : bool")]
    scope_block = [t.AnyType("scope : object"), t.AnyType("block : object")]
    sps = {
        # inference_sps.py:234-241 (resimulation_mh / mh), :253-267 (gibbs): transition_oper_type()
        "resimulation_mh": inf.sequenced_sp(transition("resimulation_mh"), inf.transition_oper_type()),
        "mh": inf.sequenced_sp(transition("resimulation_mh"), inf.transition_oper_type()),
        "gibbs": inf.sequenced_sp(transition("gibbs"), inf.transition_oper_type()),
        "emap": inf.sequenced_sp(transition("emap"), inf.par_transition_oper_type()),     # :269-280
        "gibbs_update": inf.sequenced_sp(transition("gibbs_update"), inf.par_transition_oper_type()),      # :729-732
        "pgibbs_update": inf.sequenced_sp(transition("pgibbs_update"), inf.par_transition_oper_type([t.IntegerType("particles : int")])),
        # :500-516 (slice), :269-300 (pgibbs / func_pgibbs)
        "slice": inf.sequenced_sp(transition("slice"), inf.transition_oper_type([t.NumberType("w : number"), t.IntegerType("m : int")])),
        "slice_doubling": inf.sequenced_sp(transition("slice_doubling"), inf.transition_oper_type([t.NumberType("w : number"), t.IntegerType("p : int")])),
        "pgibbs": inf.sequenced_sp(transition("pgibbs"), inf.par_transition_oper_type([t.IntegerType("particles : int")])),
        "func_pgibbs": inf.sequenced_sp(transition("func_pgibbs"), inf.par_transition_oper_type([t.IntegerType("particles : int")])),
        # :678-694
        "subsampled_mh": inf.sequenced_sp(subsampled, inf.transition_oper_type(subsampled_extra)),
        # :696-720
        "subsampled_mh_check_applicability": inf.sequenced_sp(transition("subsampled_mh_check_applicability"), inf.transition_oper_type()),
        "subsampled_mh_make_consistent": inf.sequenced_sp(transition("subsampled_mh_make_consistent"), inf.transition_oper_type(subsampled_extra[3:])),
        # :722-727
        "mh_kernel_update": inf.sequenced_sp(transition("mh_kernel_update"), inf.transition_oper_type(subsampled_extra[3:])),
        # :536-551
        "resample": inf.sequenced_sp(resample, inf.infer_action_maker_type([t.IntegerType("particles : int")])),
        # :602-606
        "likelihood_weight": inf.sequenced_sp(likelihood_weight, inf.infer_action_maker_type([])),
        "incorporate": inf.sequenced_sp(incorporate, inf.infer_action_maker_type([], return_type=t.Array(t.Number))),
        "log_likelihood_at": inf.sequenced_sp(ll_at, inf.infer_action_maker_type(scope_block, return_type=t.Array(t.Number), min_req_args=1)),
        "log_joint_at": inf.sequenced_sp(lj_at, inf.infer_action_maker_type(scope_block, return_type=t.Array(t.Number), min_req_args=1)),
    }
    return sps


def __venture_start__(ripl, model=None, n_particles=1, seed=1, device=0, trace_set=None, override_builtins=False):
    """Entry point of ripl.load_plugin.  Returns the plugin object so the caller
    can observe data through `plugin.trace_set.observe(...)`."""
    if model is None and trace_set is None:
        raise VentureException("invalid_argument", "the b200 plugin needs a compiled model (venturecxx_b200.models)")
    plugin = B200Plugin(model, n_particles, seed, device, trace_set)
    if hasattr(ripl, "bind_foreign_inference_sp"):
        for name, sp in inference_sps(plugin).items():
            ripl.bind_foreign_inference_sp(PREFIX + name, sp)
            if override_builtins:
                ripl.bind_foreign_inference_sp(name, sp)
    for name, fn in plugin.callbacks().items():
        ripl.bind_callback(name, fn)
    return plugin
