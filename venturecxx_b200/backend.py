"""The reference's own engine over the B200 particle path: backend object, backend trace, Master, TraceSet.

The reference plugs a model backend in at three seams:

  B1  a `Backend` object with `trace_constructor()` and `name()`            python/lib/shortcuts/__init__.py:90-149
  B2  the backend trace whose methods `engine.trace.Trace` forwards to      python/lib/engine/trace.py:43-126
      and the Master that fans `TraceSet` commands out over the traces      python/lib/multiprocess.py:122-271
      (`map`, `map_chunk`, `at`, `at_distinguished`, `retrieve`, `retrieve_all`, `reset_seeds`)
  --  `TraceSet` itself, whose hot methods are `self.traces.map(<method>)`  python/lib/engine/trace_set.py:148-164,
      and whose `resample` copies traces one by one                          202-235, 287-294, 339-340, 370-375

Here ONE `B200Trace` object stands for ALL particles of the set (a device-resident batch behind the C ABI,
trace_set.B200TraceSet), `B200Master` answers every Master command with one call on that batch and hands the
per-particle results back as the list `MasterBase.map` returns, and `B200EngineTraceSet` (a subclass of the
HOST Venture's TraceSet, made on first use) overrides `_trace_master`, `_resample_traces`, `copy_trace`,
`resample`, `incorporate`, `likelihood_weight` and the `log_weights` attribute so that weights live on the
device and `resample` is the resampling kernel.  The reference `Engine`, its `Infer` object, the inference
SPs of backend/lite/inference_sps.py and inference programs such as

    (do (resample 1000) (likelihood_weight) (resimulation_mh default one 160))

run UNCHANGED: `B200(model).make_engine(seed)` (or `install(engine)` on an existing engine).

Model directives.  `B200()` COMPILES the program from the `assume` / `observe` directives the engine evaluates
(compiler.py: any fixed-structure program over the scalar SPs; CompilingB200Trace).  `B200(model)` with a
hand-described model of models.py (the kernels written for C1-C5) CHECKS the directives against the program
that model stands for (`Model.program`) and decodes the observed values from the observe expressions; a
directive that does not belong to it raises VentureException("invalid_argument"), and bulk data can be given
through `engine.model.b200.observe(...)` (B200TraceSet.observe) instead of one directive per datum.

This module imports the host's `venture` package lazily (only the engine / TraceSet classes); nothing here
touches the oracle.
"""
import random

import numpy as np

from . import interchange
from .exception import VentureException
from .trace_set import B200TraceSet


class PerParticle(list):
    """A result with one entry per particle: what MasterBase.map returns for it."""


def _unsupported(what):
    return VentureException("evaluation", "the b200 backend does not support %s (per-trace tools; SURVEY 8(e) E4)" % (what,))


# ---- expressions ------------------------------------------------------------------------------------
def _norm(e, rename=None):
    """stack-dict expression (python/lib/value/dicts.py) or VentureValue -> hashable structure; numbers as
    floats, symbols through `rename` (actual name -> the compiled program's name)"""
    if hasattr(e, "asStackDict"):
        e = e.asStackDict()
    if isinstance(e, dict):
        t, v = e.get("type"), e.get("value")
        if t == "symbol":
            return ("sym", rename.get(v, v) if rename else v)
        if t in ("number", "integer", "real", "count", "probability", "atom", "positive"):
            return float(v)
        if t == "boolean":
            return bool(v)
        if isinstance(v, (list, tuple)):
            return (t, tuple(_norm(x, rename) for x in v))
        return (t, v if not isinstance(v, dict) else _norm(v, rename))
    if isinstance(e, (list, tuple)):
        return tuple(_norm(x, rename) for x in e)
    if isinstance(e, bool):
        return e
    if isinstance(e, (int, float, np.integer, np.floating)):
        return float(e)
    if isinstance(e, str):
        return ("sym", rename.get(e, e) if rename else e)
    return e


def _value(v):
    """observed value (stack dict, VentureValue or python scalar) -> python scalar"""
    if hasattr(v, "asStackDict"):
        v = v.asStackDict()
    return interchange.unbox(v)


def _is_app(n, op, n_args=None):
    return isinstance(n, tuple) and len(n) >= 1 and n[0] == ("sym", op) and (n_args is None or len(n) == n_args + 1)


class DirectiveDecoder(object):
    """Checks `assume`s against the compiled model's program and turns `observe`s into B200TraceSet.observe
    arguments.  One subclass per model kind (the models of models.py)."""

    def __init__(self, model):
        self.model = model
        self.expected = [d for d in model.program([]) if d[0] == "define"]
        self.n_defined = 0
        self.rename = {}          # the user's symbol -> the compiled program's symbol

    def define(self, name, exp):
        """-> index of the define in the compiled program"""
        if self.n_defined >= len(self.expected):
            raise VentureException("invalid_argument", "the compiled %s has no further assume directives (got %r)"
                                   % (type(self.model).__name__, name))
        _, want_name, want_exp = self.expected[self.n_defined]
        if _norm(exp, self.rename) != _norm(want_exp):
            raise VentureException("invalid_argument", "assume %r does not match directive %d of the compiled %s: expected %r"
                                   % (name, self.n_defined + 1, type(self.model).__name__, want_exp))
        self.rename[name] = want_name
        self.n_defined += 1
        return self.n_defined - 1

    def complete(self):
        return self.n_defined == len(self.expected)

    def observe(self, exp, value, n_obs):
        raise VentureException("invalid_argument", "observe directives are not decoded for %s; use engine.model.b200.observe(...)"
                               % (type(self.model).__name__,))

    def site_of_define(self, index):
        """site id whose value an `assume` reports, or None for deterministic assumes"""
        return None


class NormalNormalDecoder(DirectiveDecoder):
    def observe(self, exp, value, n_obs):
        n = _norm(exp, self.rename)
        m = self.model
        if not (_is_app(n, "normal", 2) and n[1] == ("sym", "x") and n[2] == m.obs_sigma):
            raise VentureException("invalid_argument", "observe expression is not (normal x %g)" % (m.obs_sigma,))
        return (float(_value(value)),)

    def site_of_define(self, index):
        return 0


class LogisticRegressionDecoder(DirectiveDecoder):
    def observe(self, exp, value, n_obs):
        n = _norm(exp, self.rename)
        ok = _is_app(n, "log_odds_flip", 1) and _is_app(n[1], "vector_dot", 2) and n[1][1] == ("sym", "w") and \
            _is_app(n[1][2], "array", self.model.D) and all(isinstance(x, float) for x in n[1][2][1:])
        if not ok:
            raise VentureException("invalid_argument", "observe expression is not (log_odds_flip (vector_dot w (array x_1 ... x_%d)))"
                                   % (self.model.D,))
        return (np.array(n[1][2][1:], dtype=np.float64), np.array([bool(_value(value))]))

    def site_of_define(self, index):
        return index if index < self.model.D else None


class HMMDecoder(DirectiveDecoder):
    def observe(self, exp, value, n_obs):
        n = _norm(exp, self.rename)
        ok = _is_app(n, "emission", 1) and _is_app(n[1], "state", 1) and n[1][1] == float(n_obs + 1)
        if not ok:
            raise VentureException("invalid_argument", "observe expression is not (emission (state %d)): times are observed in order"
                                   % (n_obs + 1,))
        return (int(_value(value)),)


class LDADecoder(DirectiveDecoder):
    def __init__(self, model):
        DirectiveDecoder.__init__(self, model)
        self.next_pos = {}

    def observe(self, exp, value, n_obs):
        n = _norm(exp, self.rename)
        if not (_is_app(n, "word", 2) and isinstance(n[1], float) and isinstance(n[2], float)):
            raise VentureException("invalid_argument", "observe expression is not (word doc_id pos)")
        d, pos = int(n[1]), int(n[2])
        if pos != self.next_pos.get(d, 0):
            raise VentureException("invalid_argument", "tokens of a document are observed in order: expected (word %d %d)"
                                   % (d, self.next_pos.get(d, 0)))
        self.next_pos[d] = pos + 1
        return (d, int(_value(value)))

    def site_of_define(self, index):
        return {2: 0, 3: 1}.get(index)     # alpha_document_topic, alpha_topic_word


def decoder_for(model):
    return {"NormalNormalModel": NormalNormalDecoder, "LogisticRegressionModel": LogisticRegressionDecoder,
            "HMMModel": HMMDecoder, "LDAModel": LDADecoder}.get(type(model).__name__, DirectiveDecoder)(model)


# ---- B2: the backend trace ----------------------------------------------------------------------------
class B200Trace(object):
    """What `Backend.trace_constructor()(seed)` returns: the method surface engine.trace.Trace forwards to
    (backend/lite/trace.py; Puma's python/lib/../new_cxx/trace.py:53-137), over ALL particles at once.
    Methods the reference calls per trace return a PerParticle list where each trace would answer for itself."""

    def __init__(self, model, seed, device=0, precision="fp64", trace_set=None):
        self.model = model
        self._ts = trace_set if trace_set is not None else B200TraceSet(model, 1, seed=seed, device=device, precision=precision)
        self.decoder = decoder_for(model)
        self._exps = {}            # did -> expression handed to eval
        self._kind = {}            # did -> ("define", index) | ("observe",) | ("evaluate",)
        self._bound = {}           # symbol -> did
        self._profiling = False

    @property
    def ts(self):
        return self._ts

    @property
    def n_particles(self):
        return self.ts.n_particles

    # -- what the TraceSet needs besides the per-trace surface ------------------------------------------
    def get_log_weights(self):
        return self.ts.log_weights.tolist()

    def set_log_weights(self, values):
        values = list(values)
        if len(values) != self.n_particles:
            raise VentureException("invalid_argument", "%d log weights for %d particles" % (len(values), self.n_particles))
        self.ts.set_particle_log_weights(np.asarray(values, dtype=np.float64))

    def resample(self, P, mode):
        self.ts.resample(int(P), mode=mode)

    # -- model building (engine/trace.py:53-126 calls these) --------------------------------------------
    def eval(self, did, exp):
        self._exps[did] = exp
        self._kind[did] = ("evaluate",)

    def bindInGlobalEnv(self, sym, did):
        index = self.decoder.define(sym, self._exps[did])
        self._kind[did] = ("define", index)
        self._bound[sym] = did

    def boundInGlobalEnv(self, sym):
        return sym in self._bound

    def unbindInGlobalEnv(self, sym):
        raise _unsupported("forgetting an assume of a compiled model")

    def uneval(self, did):
        kind = self._kind.pop(did, None)
        self._exps.pop(did, None)
        if kind is not None and kind[0] == "define":
            raise _unsupported("forgetting an assume of a compiled model")

    def observe(self, did, val):
        args = self.decoder.observe(self._exps[did], val, len(self.ts.observations()))
        self.ts.observe(*args)
        self._kind[did] = ("observe",)

    def unobserve(self, did):
        """engine.forget of an observation.  The device drops ALL observations at once
        (vk_forget_observations), so only forgetting the whole data set -- one directive after the other, most
        recent first is not required -- is expressible: the weights are reported with the first forget."""
        if self._kind.get(did, (None,))[0] != "observe":
            raise VentureException("invalid_argument", "directive %r is not an observation" % (did,))
        obs = [d for d, k in self._kind.items() if k[0] == "observe"]
        if len(self.ts.observations()) > 0:
            if len(obs) < len(self.ts.observations()):
                raise _unsupported("forgetting single observations next to bulk-observed data")
            w = self.ts.forget_observations()
            if isinstance(self.decoder, LDADecoder):
                self.decoder.next_pos = {}
            return PerParticle(w.tolist())
        return PerParticle([0.0] * self.n_particles)

    def bindPrimitiveSP(self, name, sp):
        raise _unsupported("foreign SPs inside a compiled model")

    def bindPrimitiveName(self, name, val):
        raise _unsupported("foreign values inside a compiled model")

    def sealEnvironment(self):
        pass

    def freeze(self, did):
        raise _unsupported("freeze")

    def _values_of(self, did):
        kind = self._kind.get(did)
        if kind is None:
            raise VentureException("invalid_argument", "no directive %r" % (did,))
        if kind[0] == "define":
            site = self.decoder.site_of_define(kind[1])
            if site is not None:
                return [interchange.number(v) for v in self.ts.read_site(site)]
            return [{"type": "string", "value": "<compiled: %s>" % (self.decoder.expected[kind[1]][1],)}] * self.n_particles
        if kind[0] == "evaluate":
            n = _norm(self._exps[did], self.decoder.rename)
            if isinstance(n, tuple) and len(n) == 2 and n[0] == "sym":
                did2 = self._bound.get(dict((v, k) for k, v in self.decoder.rename.items()).get(n[1], n[1]))
                if did2 is not None:
                    return self._values_of(did2)
            if isinstance(n, float):
                return [interchange.number(n)] * self.n_particles
            raise _unsupported("predicting %r on a compiled model" % (self._exps[did],))
        raise _unsupported("reading the value of an observation")

    def extractValue(self, did):
        return self._values_of(did)[0]

    def extractValues(self, did):
        return PerParticle(self._values_of(did))

    def extractRaw(self, did):
        return self.extractValue(did)

    # -- the hot path ------------------------------------------------------------------------------------
    def makeConsistent(self):
        return PerParticle(self.ts.incorporate().tolist())

    def registerConstraints(self):
        self.ts.incorporate(return_increments=False)

    def likelihood_weight(self):
        self.ts.likelihood_weight()
        return PerParticle(self.ts.log_weights.tolist())

    def primitive_infer(self, exp):
        """exp = [operator, scope, block, ...extra..., transitions] of VentureValues / stack dicts / python values
        (backend/lite/infer/dispatch.py:50-107)."""
        args = [_infer_arg(e) for e in exp]
        out = np.asarray(self.ts.primitive_infer(args))
        if args and args[0] == "subsampled_mh" and getattr(self.ts, "last_mean_nodes", None) is not None:
            out = np.asarray(self.ts.last_mean_nodes)       # the engine gets what Lite returns: node counts, not the test statistics
        return PerParticle(out[..., None].reshape(self.n_particles, -1)[:, 0].tolist())

    def log_likelihood_at(self, scope, block):
        return PerParticle(self.ts.log_likelihood_at(_infer_arg(scope), _infer_arg(block)).tolist())

    def log_joint_at(self, scope, block):
        return PerParticle(self.ts.log_joint_at(_infer_arg(scope), _infer_arg(block)).tolist())

    def reset_to_prior(self):
        self.ts.reset_to_prior()

    def numRandomChoices(self):
        return int(self.ts.num_sites())

    def numBlocksInScope(self, scope):
        return int(self.ts.numBlocksInScope(_infer_arg(scope)))

    def scope_keys(self):
        return sorted(getattr(self.model, "scopes", {"default": 0}))

    # -- copy / serialise ----------------------------------------------------------------------------------
    def short_circuit_copyable(self):
        return False

    def stop_and_copy(self):
        raise _unsupported("copying single traces (resample runs on the device)")

    def dump(self, skipStackDictConversion=False):
        """per particle, what engine.trace.Trace.dump returns for a Lite trace holding it (interchange.py)"""
        return PerParticle(self.ts.dump())

    # -- RNG -------------------------------------------------------------------------------------------------
    def has_own_prng(self):
        return True

    def set_seed(self, seed):
        self.ts.set_seed(int(seed))

    def get_seed(self):
        return self.ts.get_seed()

    # -- misc ------------------------------------------------------------------------------------------------
    def set_profiling(self, enabled=True):
        self._profiling = bool(enabled)

    def clear_profiling(self):
        pass

    def stop(self):
        pass

    def close(self):
        self.ts.close()

    def __getattr__(self, name):
        if name in ("select", "just_detach", "just_regen", "just_restore", "detach_for_proposal", "regen_with_proposal",
                    "get_current_values", "block_values", "diversify", "checkInvariants", "makeSerializationDB",
                    "dumpSerializationDB", "unevalAndExtract", "restore", "evalAndRestore"):
            def refuse(*_a, **_k):
                raise _unsupported(name)
            return refuse
        raise AttributeError(name)


class CompilingB200Trace(B200Trace):
    """B200Trace without a hand-described model: the `assume` / `observe` directives the engine evaluates are
    RECORDED and compiled (compiler.py) when the particle set is first needed; directives that
    arrive later extend the program -- the existing sites keep their values, new sites are simulated from their
    priors given them (vk_program_resimulate), and new observations contribute their log density to the weights
    exactly as Trace.makeConsistent does for a leaf observation (lite/trace.py:422-448)."""

    def __init__(self, seed, device=0, precision="fp64"):
        self.model = None
        self._seed, self._device, self._precision = int(seed), int(device), precision
        self._ts, self._dirty = None, False
        self.source = []                      # ("assume", name, exp) | ("observe", exp, value), in directive order
        self._pending_incr = None
        self._exps, self._kind, self._bound = {}, {}, {}
        self._define_site = {}                # did -> site index, for assumes whose root is a random application
        self._profiling = False
        self.decoder = None
        self._n_initial, self._initial_lw = 1, None

    @property
    def ts(self):
        self._materialize()
        return self._ts

    @property
    def n_particles(self):
        return self._ts.n_particles if self._ts is not None else self._n_initial

    def ready(self):
        """a particle set exists once the program has a random choice"""
        return self._ts is not None or bool(getattr(self, "_names", None))

    def get_log_weights(self):
        if not self.ready():
            return list(self._initial_lw) if self._initial_lw is not None else [0.0] * self._n_initial
        return B200Trace.get_log_weights(self)

    def set_log_weights(self, values):
        if not self.ready():
            values = list(values)
            if len(values) != self._n_initial:
                raise VentureException("invalid_argument", "%d log weights for %d particles" % (len(values), self._n_initial))
            self._initial_lw = [float(v) for v in values]
            return
        B200Trace.set_log_weights(self, values)

    def resample(self, P, mode):
        if not self.ready():                   # nothing random yet: P identical empty traces
            lw = self.get_log_weights()
            m = max(lw)
            lse = m if m == float("-inf") else m + float(np.log(np.sum(np.exp(np.asarray(lw) - m))))
            self._n_initial = int(P)
            self._initial_lw = [lse - float(np.log(P))] * int(P)
            return
        B200Trace.resample(self, P, mode)

    def _materialize(self):
        if self._ts is not None and not self._dirty:
            return
        from . import _lib
        from .compiler import CompiledModel
        model = CompiledModel(self.source)
        old = self._ts
        P = old.n_particles if old is not None else self._n_initial
        new = B200TraceSet(model, P, seed=self._seed, device=self._device, precision=self._precision)
        if old is None:
            if self._initial_lw is not None:
                new.set_particle_log_weights(np.asarray(self._initial_lw))
            incr = new.incorporate()
        else:
            n_old = old.num_sites()
            step = old.rng_step
            for j in range(n_old):
                new.write_site(j, old.read_site(j))
            new.rng_step = step
            _lib.call("vk_program_resimulate", new._h, n_old)
            lw_old = old.log_weights
            ll_old = old.log_likelihood_at("default", "all")
            new.incorporate(return_increments=False)
            incr = new.log_likelihood_at("default", "all") - ll_old
            incr = np.where(np.isnan(incr), -np.inf, incr)
            new.set_particle_log_weights(lw_old + incr)
            old.close()
        self._ts, self.model, self._dirty = new, model, False
        self._pending_incr = incr if self._pending_incr is None else self._pending_incr + incr

    # -- directives ------------------------------------------------------------------------------------------
    def bindInGlobalEnv(self, sym, did):
        self.source.append(("assume", sym, self._exps[did]))
        self._check()
        self._kind[did] = ("define", len(self.source) - 1)
        self._bound[sym] = did
        self._dirty = True

    def observe(self, did, val):
        self.source.append(("observe", self._exps[did], val))
        self._check()
        self._kind[did] = ("observe",)
        self._dirty = True

    def _check(self):
        """compile errors belong to the directive that causes them (the engine reports them to the user)"""
        from .compiler import Compiler
        c = Compiler()
        try:
            for d in self.source:
                if d[0] == "assume":
                    c.assume(d[1], d[2])
                else:
                    c.observe(d[1], d[2])
        except VentureException:
            self.source.pop()
            raise
        self._names = list(c.site_names)

    def unobserve(self, did):
        raise _unsupported("forgetting an observation of a compiled program")

    def _values_of(self, did):
        kind = self._kind.get(did)
        if kind is None:
            raise VentureException("invalid_argument", "no directive %r" % (did,))
        if kind[0] == "define":
            name = self.source[kind[1]][1]
            ts = self.ts
            if name in self.model.site_names:
                j = self.model.site_names.index(name)
                c = self.model.compiler
                boolean = c.nodes[c.sites[j]].op in ("FLIP", "LOG_FLIP", "LOG_ODDS_FLIP")
                return [interchange.boolean(v != 0.0) if boolean else interchange.number(v) for v in ts.read_site(j)]
            return [{"type": "string", "value": "<compiled: %s>" % (name,)}] * self.n_particles
        if kind[0] == "evaluate":
            from .compiler import _plain
            e = _plain(self._exps[did])
            if isinstance(e, str) and e in self._bound:
                return self._values_of(self._bound[e])
            if isinstance(e, float):
                return [interchange.number(e)] * self.n_particles
            raise _unsupported("predicting %r on a compiled program" % (self._exps[did],))
        raise _unsupported("reading the value of an observation")

    def makeConsistent(self):
        if not self.ready():
            return PerParticle([0.0] * self._n_initial)
        self._materialize()
        incr, self._pending_incr = self._pending_incr, None
        if incr is None:
            incr = np.zeros(self.n_particles)
        return PerParticle(np.asarray(incr).tolist())

    def scope_keys(self):
        return ["default"]

    def set_seed(self, seed):
        self._seed = int(seed)
        if self._ts is not None:
            self._ts.set_seed(int(seed))

    def get_seed(self):
        return self._ts.get_seed() if self._ts is not None else self._seed

    def numRandomChoices(self):
        return int(self.ts.num_sites()) if self.ready() else 0

    def reset_to_prior(self):
        if self.ready():
            self.ts.reset_to_prior()

    def close(self):
        if self._ts is not None:
            self._ts.close()


def _infer_arg(v):
    """argument of an inference operator: VentureValue / stack dict / python value -> python value"""
    if hasattr(v, "asStackDict"):
        v = v.asStackDict()
    if isinstance(v, dict) and "value" in v:
        v = v["value"]
    if isinstance(v, float) and v == int(v):
        return int(v)
    return v


# ---- B2: the Master -------------------------------------------------------------------------------------
class B200Master(object):
    """python/lib/multiprocess.py:122-271 (MasterBase) over ONE batch: `map` is one call on the device batch."""

    _DIRECT = ("makeConsistent", "likelihood_weight", "primitive_infer", "log_likelihood_at", "log_joint_at",
               "reset_to_prior", "dump", "numRandomChoices", "get_seed", "set_seed", "set_profiling", "clear_profiling")

    def __init__(self, objects, process_cap=None, seed=None):
        if len(objects) != 1:
            raise _unsupported("a pool of %d separate traces (one batch object holds every particle)" % (len(objects),))
        self.wrapper = objects[0]                       # engine.trace.Trace around the B200Trace (directive records)
        self.batch = getattr(objects[0], "trace", objects[0])
        self.process_cap = process_cap
        if seed is not None:
            self.reset_seeds(seed)

    def __del__(self):
        pass

    def _n(self):
        return self.batch.n_particles

    def reset_seeds(self, seed):
        # multiprocess.py:186-191 draws one seed per trace from Random(seed); the batch derives per-particle
        # Philox streams from ONE seed, which it already got from trace_constructor()(seed)
        pass

    def _call(self, cmd, args, kwargs):
        target = getattr(self.batch, cmd) if cmd in self._DIRECT else getattr(self.wrapper, cmd)
        return target(*args, **kwargs)

    def map(self, cmd, *args, **kwargs):
        if cmd == "stop":
            return None
        if cmd == "evaluate":
            self.wrapper.evaluate(*args, **kwargs)
            return list(self.batch.extractValues(args[0]))
        if cmd == "send_object":
            return self.retrieve_all()
        res = self._call(cmd, args, kwargs)
        return list(res) if isinstance(res, PerParticle) else [res] * self._n()

    def map_chunk(self, ix, cmd, *args, **kwargs):
        if ix != 0:
            raise IndexError("the b200 master has one chunk")
        if cmd == "set_seeds":
            return None
        return self.map(cmd, *args, **kwargs)

    def at(self, ix, cmd, *args, **kwargs):
        if not 0 <= ix < self._n():
            raise IndexError("particle %d of %d" % (ix, self._n()))
        if cmd == "send_object":
            return self.retrieve(ix)
        if cmd == "dump":
            return self.batch.ts.dump_particle(ix)
        if cmd in ("report_value", "report_raw"):
            self.wrapper.directive(args[0])            # raises for an unknown directive, as the reference does
            return self.batch._values_of(args[0])[ix]
        res = self._call(cmd, args, kwargs)
        return res[ix] if isinstance(res, PerParticle) else res

    def at_distinguished(self, cmd, *args, **kwargs):
        return self.at(0, cmd, *args, **kwargs)

    def can_shortcut_retrieval(self):
        return True

    def retrieve(self, ix):
        """particle ix as a real Lite trace of the host Venture (engine/trace.py:168-171): the device state,
        restored through the reference's own serialisation format (interchange.py)"""
        import venture.engine.trace as tr
        from venture.lite.trace import Trace as LiteTrace
        rng = random.Random(self.batch.get_seed() + ix)
        return tr.Trace.restore(lambda: LiteTrace(rng.randint(1, 2 ** 31 - 1)), self.batch.ts.dump_particle(ix), {})

    def retrieve_all(self):
        return [self.retrieve(i) for i in range(self._n())]


# ---- the TraceSet ---------------------------------------------------------------------------------------
_trace_set_class = None


def trace_set_class():
    """B200EngineTraceSet: a subclass of the HOST Venture's TraceSet (python/lib/engine/trace_set.py)."""
    global _trace_set_class
    if _trace_set_class is not None:
        return _trace_set_class
    from venture.engine.trace_set import TraceSet

    class B200EngineTraceSet(TraceSet):
        """trace_set.py:36-414 with the particle set on the device."""

        def _trace_master(self, mode):            # trace_set.py:50-60
            return B200Master

        @property
        def b200(self):
            return self.traces.batch.ts

        # weights live on the device; the reference reads and assigns `log_weights` as a list attribute
        @property
        def log_weights(self):
            if getattr(self, "traces", None) is None:
                return []
            return self.traces.batch.get_log_weights()

        @log_weights.setter
        def log_weights(self, values):
            self.traces.batch.set_log_weights(values)

        def likelihood_weight(self):              # :287-288 (the kernel overwrites the device weights itself)
            self.traces.map("likelihood_weight")

        def incorporate(self):                    # :290-294 (the kernel adds the increments itself)
            return self.traces.map("makeConsistent")

        def resample(self, P, mode="sequential", process_cap=None):      # :202-207
            self.mode = mode
            self.process_cap = process_cap
            self._resample_traces(P)
            self.incorporate()

        def _resample_traces(self, P):            # :209-219 -> the resampling kernel; weights evened out on the device
            self.traces.batch.resample(int(P), getattr(self.backend, "resample_mode", "multinomial"))
            return None

        def copy_trace(self, trace):              # :370-375
            raise _unsupported("copying single traces (resample runs on the device)")

        def reinit_inference_problem(self, num_particles=1):             # :187-200
            self.resample(num_particles)
            self.log_weights = [0.0] * int(num_particles)
            self.traces.map("reset_to_prior")

        def diversify(self, program):
            raise _unsupported("diversify")

        def _collapse_help(self, scope, block, select_keeper):
            raise _unsupported("collapse_equal / collapse_equal_map")

        def for_each_trace_sequential(self, f):
            raise _unsupported("for_each_particle")

        def on_trace(self, i, f):
            raise _unsupported("on_particle")

    _trace_set_class = B200EngineTraceSet
    return _trace_set_class


def install(engine, backend=None, seed=None):
    """Swap the B200 trace set into a reference Engine (engine.py:35-38, 432-435): `engine.model` becomes a
    B200EngineTraceSet over `backend` and `engine.new_model` makes further ones."""
    import types
    cls = trace_set_class()
    backend = backend if backend is not None else engine.model.backend
    old = engine.model

    def new_model(self, backend_=None):
        b = backend_ if backend_ is not None else self.model.backend
        s = self._py_rng.randint(1, 2 ** 31 - 1)
        if isinstance(b, B200):
            return cls(self, b, s)
        from venture.engine.trace_set import TraceSet
        return TraceSet(self, b, s)

    engine.new_model = types.MethodType(new_model, engine)
    engine.model = cls(engine, backend, seed if seed is not None else engine._py_rng.randint(1, 2 ** 31 - 1))
    try:
        batch = getattr(old.traces, "batch", None)
        if batch is None:
            batch = getattr(old.traces.at_distinguished("send_object"), "trace", None)
        if isinstance(batch, B200Trace):
            batch.close()
    except Exception:
        pass
    return engine


# ---- B1: the backend object --------------------------------------------------------------------------------
class B200(object):
    """python/lib/shortcuts/__init__.py:90-149 (`Backend`, `Lite`, `Puma`): `backend("b200")`.
    model: a compiled model of venturecxx_b200.models; resample_mode: "multinomial" (the reference's law,
    trace_set.py:209-219) or "systematic"."""

    def __init__(self, model=None, device=0, resample_mode="multinomial", precision="fp64"):
        """model=None: the program is COMPILED from the directives the engine evaluates (compiler.py)."""
        self.model, self.device, self.resample_mode, self.precision = model, int(device), resample_mode, precision

    def trace_constructor(self):
        if self.model is None:
            return lambda seed: CompilingB200Trace(seed, device=self.device, precision=self.precision)
        return lambda seed: B200Trace(self.model, seed, device=self.device, precision=self.precision)

    def name(self):
        return "b200"

    def make_engine(self, persistent_inference_trace=True, seed=None):
        from venture.engine.engine import Engine
        seed = seed if seed is not None else random.randint(1, 2 ** 31 - 1)
        return install(Engine(self, seed, persistent_inference_trace), self)

    def _host(self):
        import venture.shortcuts as s
        return s

    def make_core_sivm(self, persistent_inference_trace=True, seed=None):
        from venture import sivm
        return sivm.CoreSivm(self.make_engine(persistent_inference_trace, seed))

    def make_venture_sivm(self, persistent_inference_trace=True, seed=None):
        from venture import sivm
        return sivm.VentureSivm(self.make_core_sivm(persistent_inference_trace, seed))

    def make_combined_ripl(self, persistent_inference_trace=True, seed=None, **kwargs):
        from venture import parser, ripl
        v = self.make_venture_sivm(persistent_inference_trace, seed)
        modes = {"church_prime": parser.ChurchPrimeParser.instance(), "venture_script": parser.VentureScriptParser.instance()}
        r = ripl.Ripl(v, modes, **kwargs)
        r.set_mode("church_prime")
        r.backend_name = self.name()
        return r

    def make_ripl(self, init_mode="venture_script", **kwargs):
        r = self.make_combined_ripl(**kwargs)
        r.set_mode(init_mode)
        return r

    def make_church_prime_ripl(self, **kwargs):
        return self.make_ripl(init_mode="church_prime", **kwargs)

    def make_venture_script_ripl(self, **kwargs):
        return self.make_ripl(init_mode="venture_script", **kwargs)


def backend(name="b200", model=None, **kwargs):
    """`venture.shortcuts.backend(name)` with the extra name "b200" (model: a hand-described model of
    venturecxx_b200.models, or None to compile the program from its directives)."""
    if name == "b200":
        return B200(model, **kwargs)
    import venture.shortcuts as s
    return s.backend(name)
