"""B200TraceSet: the device-resident stand-in for the reference's TraceSet
(python/lib/engine/trace_set.py:36-414) over N traces of one compiled model.

Method names, argument meaning and return shapes follow the reference:
  likelihood_weight   trace_set.py:287-288   (overwrites log_weights)
  incorporate         trace_set.py:290-294   (adds makeConsistent increments)
  resample            trace_set.py:202-207   (+ systematic mode)
  primitive_infer     trace_set.py:339-340 -> backend/lite/infer/dispatch.py:130-306
  log_likelihood_at / log_joint_at           dispatch.py:331-361
  reset_to_prior      trace_set.py:199-200
Where the reference returns a list with one entry per trace (traces.map), a
numpy array of length P is returned.
"""
import ctypes

import numpy as np

from . import _lib
from .exception import VentureException

_BLOCKS = {"one": _lib.BLOCK_ONE, "all": _lib.BLOCK_ALL, "each": _lib.BLOCK_EACH, "ordered": _lib.BLOCK_ORDERED}


def _unbox(v):
    """Accept stack dicts {"type":..,"value":..} (python/lib/value/dicts.py) or raw values."""
    if isinstance(v, dict) and "value" in v:
        return v["value"]
    return v


class B200TraceSet(object):
    def __init__(self, model, n_particles, seed=1, device=0, particle_base=0, precision="fp64"):
        self.model = model
        self.device = int(device)
        self._h = model._create(int(n_particles), int(seed), self.device)
        self.n_particles = int(n_particles)
        self._observed = []
        self.last_mean_nodes = None
        if particle_base:
            _lib.call("vk_set_particle_base", self._h, int(particle_base))
        # `assume` simulates every latent from its prior (regen.py eval)
        _lib.call("vk_reset_to_prior", self._h)
        if precision != "fp64":
            self.precision = precision

    # -- lifetime --------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            _lib.load().vk_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def use_stream(self, cuda_stream):
        """Run on a caller stream (e.g. torch.cuda.current_stream().cuda_stream)."""
        _lib.call("vk_set_stream", self._h, ctypes.c_void_p(cuda_stream))

    def synchronize(self):
        _lib.call("vk_synchronize", self._h)

    # -- precision of the hot loop (include/venture_b200.h vk_set_precision) --------
    @property
    def precision(self):
        v = ctypes.c_int32()
        _lib.call("vk_get_precision", self._h, ctypes.byref(v))
        return {_lib.FP64: "fp64", _lib.FP32: "fp32"}[v.value]

    @precision.setter
    def precision(self, name):
        if name not in ("fp64", "fp32"):
            raise VentureException("invalid_argument", "precision must be 'fp64' or 'fp32', not %r" % (name,))
        _lib.call("vk_set_precision", self._h, _lib.FP32 if name == "fp32" else _lib.FP64)

    # -- seeds (trace_set.py has_own_prng/set_seed) ------------------------
    def set_seed(self, seed):
        _lib.call("vk_set_seed", self._h, int(seed))

    def get_seed(self):
        s = ctypes.c_uint64()
        _lib.call("vk_get_seed", self._h, ctypes.byref(s))
        return s.value

    @property
    def rng_step(self):
        s = ctypes.c_uint32()
        _lib.call("vk_get_step", self._h, ctypes.byref(s))
        return s.value

    @rng_step.setter
    def rng_step(self, v):
        _lib.call("vk_set_step", self._h, int(v))

    # -- model building ----------------------------------------------------
    def observe(self, *args):
        """engine.observe -> TraceSet.observe (trace_set.py:154-155): values stay
        unpropagated until incorporate()."""
        self.model._observe(self._h, *args)
        self._observed.append(args)       # host record of the observe directives (interchange.py); not copied

    def observations(self):
        """The observe calls made so far, in order, as given (argument tuples)."""
        return list(self._observed)

    # -- state interchange with Lite (interchange.py) ----------
    def dump_particle(self, index):
        """What engine.trace.Trace.dump() returns for a Lite trace holding particle `index`."""
        from . import interchange
        return interchange.dump_particle(self, int(index))

    def dump(self):
        from . import interchange
        return interchange.dump(self)

    def load(self, serialized):
        from . import interchange
        interchange.load(self, serialized)

    def forget_observations(self, return_weights=True):
        """engine.forget of every observation (trace_set.py:157-161): returns the
        per-particle weights unconstrain() reports; log_weights unchanged.
        return_weights=False skips computing them (a full pass over the observations' log densities)
        and keeps the call stream-ordered."""
        out = self._out() if return_weights else None
        _lib.call("vk_forget_observations", self._h, _lib.ptr(out))
        self._observed = []
        return out

    def reset_to_prior(self):
        _lib.call("vk_reset_to_prior", self._h)

    # -- weights -----------------------------------------------------------
    def _out(self, n=None, dtype=np.float64):
        return np.empty(self.n_particles if n is None else n, dtype=dtype)

    @property
    def log_weights(self):
        out = self._out()
        _lib.call("vk_get_log_weights", self._h, _lib.ptr(out))
        return out

    @log_weights.setter
    def log_weights(self, values):
        self.set_particle_log_weights(values)

    def particle_log_weights(self):           # engine/inference.py:229-230
        return self.log_weights

    def set_particle_log_weights(self, values):   # engine/inference.py:232-234
        v = np.ascontiguousarray(values, dtype=np.float64)
        if v.shape != (self.n_particles,):
            raise VentureException("invalid_argument", "expected %d log weights" % self.n_particles)
        _lib.call("vk_set_log_weights", self._h, _lib.ptr(v))

    def increment_particle_log_weights(self, values):
        v = np.ascontiguousarray(values, dtype=np.float64)
        if v.shape != (self.n_particles,):
            raise VentureException("invalid_argument", "expected %d log weights" % self.n_particles)
        _lib.call("vk_increment_log_weights", self._h, _lib.ptr(v))

    def equalize_particle_log_weights(self):      # utils.log_domain_even_out
        lse = self.logsumexp()
        self.set_particle_log_weights(np.full(self.n_particles, lse - np.log(self.n_particles)))

    def particle_normalized_probs(self):          # utils.logWeightsToNormalizedDirect
        out = self._out()
        _lib.call("vk_normalized_probs", self._h, _lib.ptr(out))
        return out

    def logsumexp(self):
        out = ctypes.c_double()
        _lib.call("vk_logsumexp", self._h, ctypes.byref(out))
        return out.value

    def likelihood_weight(self):
        _lib.call("vk_likelihood_weight", self._h)

    def incorporate(self, return_increments=True):
        """trace_set.py:290-294.  The reference returns the per-trace increments; pass
        return_increments=False to keep the call stream-ordered (no device-to-host copy)."""
        out = self._out() if return_increments else None
        _lib.call("vk_make_consistent", self._h, _lib.ptr(out))
        return out

    def resample(self, P=None, mode="multinomial", uniforms=None, return_ancestors=False):
        """trace_set.py:202-207.  mode: "multinomial" (reference semantics) or
        "systematic".  `uniforms` overrides the stream draws (parity tests)."""
        P = self.n_particles if P is None else int(P)
        m = {"multinomial": _lib.RESAMPLE_MULTINOMIAL, "systematic": _lib.RESAMPLE_SYSTEMATIC,
             "sequential": _lib.RESAMPLE_MULTINOMIAL}.get(mode)
        if m is None:
            raise VentureException("invalid_argument", "unknown resample mode %r" % (mode,))
        u = None if uniforms is None else np.ascontiguousarray(np.atleast_1d(uniforms), dtype=np.float64)
        anc = np.empty(P, dtype=np.int32) if return_ancestors else None
        _lib.call("vk_resample", self._h, P, m, _lib.ptr(u), _lib.ptr(anc))
        self.n_particles = P             # trace_set.py:202-219: the set now holds P particles
        return anc

    # -- inference ---------------------------------------------------------
    def _scope_block(self, scope, block):
        scope, block = _unbox(scope), _unbox(block)
        scopes = getattr(self.model, "scopes", {"default": 0})
        if scope in scopes:
            sid = scopes[scope]
        elif scope in scopes.values():
            sid = int(scope)
        else:
            raise VentureException("invalid_argument", "scope %r is not compiled for this model (have %s)"
                                   % (scope, sorted(scopes)))
        # a scope made by `tag` names its blocks by the program's own values (numbers or symbols): -> the block's index
        tagged = getattr(self.model, "scope_blocks", {}).get(sid)
        if tagged is not None:
            key = float(block) if isinstance(block, (int, float)) and not isinstance(block, bool) else block
            if key in tagged:
                return sid, tagged[key]
            if not (isinstance(block, str) and block in _BLOCKS):
                raise VentureException("invalid_argument", "scope %r has no block %r (have %s)" % (scope, block, sorted(tagged, key=str)))
        if isinstance(block, str):
            if block not in _BLOCKS:
                raise VentureException("invalid_argument", "unknown block %r" % (block,))
            return sid, _BLOCKS[block]
        return sid, int(block)

    def primitive_infer(self, exp):
        """exp = [operator, scope, block, ...extra..., transitions]
        (backend/lite/infer/dispatch.py:50-107, 130-161)."""
        exp = [_unbox(e) for e in exp]
        op = exp[0]
        if op in ("resimulation_mh", "mh"):
            if len(exp) != 4:
                raise VentureException("invalid_argument", "%s takes (scope block transitions)" % op)
            return self.resimulation_mh(exp[1], exp[2], exp[3])
        if op == "mh_kernel_update":
            # dispatch.py:145-148, 308-313: mh over a scaffold built with (useDeltaKernels deltaKernelArgs updateValues).  None of
            # the compiled SPs has a delta kernel (only MVNormalOutputPSP does, continuous.py:121-124) and no value on the
            # device is ever stale, so this is resimulation_mh
            if len(exp) != 7:
                raise VentureException("invalid_argument", "mh_kernel_update takes (scope block useDeltaKernels deltaKernelArgs updateValues transitions)")
            return self.resimulation_mh(exp[1], exp[2], exp[6])
        # gibbs_update / pgibbs_update (dispatch.py:185-189, 213-226): the same operators over scaffolds built with
        # updateValues=True, i.e. tolerating values left stale by subsampled_mh -- nothing is stale on the device
        if op == "gibbs_update":
            op = "gibbs"
        if op == "pgibbs_update":
            op = "pgibbs"
        if op in ("gibbs", "emap"):
            if len(exp) == 5 and isinstance(exp[4], bool):      # the trailing in_parallel flag (par_transition_oper_type)
                exp = exp[:4]
            if len(exp) != 4:
                raise VentureException("invalid_argument", "%s takes (scope block transitions)" % op)
            return getattr(self, op)(exp[1], exp[2], exp[3])
        if op in ("pgibbs", "func_pgibbs"):
            if len(exp) == 6 and isinstance(exp[5], bool):      # the trailing in_parallel flag (inference_sps.py:269-300)
                exp = exp[:5]
            if len(exp) != 5:
                raise VentureException("invalid_argument", "%s takes (scope block particles transitions)" % op)
            return self.pgibbs(exp[1], exp[2], exp[3], exp[4])
        if op == "slice":
            if len(exp) != 6:
                raise VentureException("invalid_argument", "slice takes (scope block w m transitions)")
            return self.slice(exp[1], exp[2], exp[3], exp[4], exp[5])
        if op == "slice_doubling":
            if len(exp) != 6:
                raise VentureException("invalid_argument", "slice_doubling takes (scope block w p transitions)")
            return self.slice_doubling(exp[1], exp[2], exp[3], exp[4], exp[5])
        if op == "subsampled_mh":
            if len(exp) != 10:
                raise VentureException(
                    "invalid_argument",
                    "subsampled_mh takes (scope block Nbatch k0 epsilon useDeltaKernels deltaKernelArgs "
                    "updateValues transitions)")
            return self.subsampled_mh(*exp[1:])
        if op == "subsampled_mh_check_applicability":
            # dispatch.py:153-157: warns, changes nothing, reports 0 nodes
            import warnings
            from . import exception as exc
            self._scope_block(exp[1], exp[2])
            check = getattr(self.model, "subsampled_applicability", None)
            if check is None:
                raise VentureException("invalid_argument", "subsampled_mh is not compiled for this model")
            ok, kind, msg = check(exp[2])
            if not ok:
                warnings.warn(msg, getattr(exc, kind))
            return np.zeros(self.n_particles)
        if op == "subsampled_mh_make_consistent":
            # SubsampledMHOperator.makeConsistent (infer/subsampled_mh.py:411-441) refreshes the local sections an accepted
            # proposal left stale.  The device keeps no per-section values (sections are evaluated from the current global
            # value; the logistic-regression model's logit cache is only MARKED stale and rebuilt by its next reader), so
            # there is nothing to do but report the node count.  (The reference's own dispatch entry, dispatch.py:158-161,
            # reads an unbound `extra` and raises NameError.)
            self._scope_block(exp[1], exp[2])
            nodes = getattr(self.model, "subsampled_global_nodes", None)
            if nodes is None:
                raise VentureException("invalid_argument", "subsampled_mh is not compiled for this model")
            return np.full(self.n_particles, nodes(exp[2]))
        raise VentureException("invalid_argument", "operator %r is not on the B200 path" % (op,))

    def resimulation_mh(self, scope, block, transitions, return_accepts=False, out=None):
        """out: optional caller-owned float64 array of n_particles (e.g. pinned memory) for the node counts"""
        s, b = self._scope_block(scope, block)
        if out is None:
            out = self._out()
        elif out.shape != (self.n_particles,) or out.dtype != np.float64 or not out.flags["C_CONTIGUOUS"]:
            raise VentureException("invalid_argument", "out must be a contiguous float64 array of %d entries" % self.n_particles)
        acc = self._out(dtype=np.int64) if return_accepts else None
        _lib.call("vk_mh", self._h, s, b, int(transitions), _lib.ptr(out), _lib.ptr(acc))
        return (out, acc) if return_accepts else out

    mh = resimulation_mh

    def gibbs(self, scope, block, transitions, return_changes=False):
        """(gibbs scope block transitions): enumerative Gibbs on discrete sites
        (infer/dispatch.py:175-179 -> infer/egibbs.py:42-107)."""
        s, b = self._scope_block(scope, block)
        out = self._out()
        ch = self._out(dtype=np.int64) if return_changes else None
        _lib.call("vk_gibbs", self._h, s, b, int(transitions), _lib.ptr(out), _lib.ptr(ch))
        return (out, ch) if return_changes else out

    def emap(self, scope, block, transitions, return_changes=False):
        """(emap scope block transitions): move to the conditional maximum found by enumeration
        (infer/dispatch.py:180-184 -> EnumerativeMAPOperator, infer/egibbs.py:117-121)."""
        s, b = self._scope_block(scope, block)
        out = self._out()
        ch = self._out(dtype=np.int64) if return_changes else None
        _lib.call("vk_emap", self._h, s, b, int(transitions), _lib.ptr(out), _lib.ptr(ch))
        return (out, ch) if return_changes else out

    def pgibbs(self, scope, block, particles, transitions, return_accepts=False):
        """(pgibbs scope ordered particles transitions): particle Gibbs over the ordered blocks of a scope
        (infer/dispatch.py:197-212 -> infer/pgibbs.py:49-126)."""
        s, b = self._scope_block(scope, block)
        out = self._out()
        acc = self._out(dtype=np.int64) if return_accepts else None
        _lib.call("vk_pgibbs", self._h, s, b, int(particles), int(transitions), _lib.ptr(out), _lib.ptr(acc))
        return (out, acc) if return_accepts else out

    def slice(self, scope, block, w, m, transitions, return_moves=False):
        """(slice scope block w m transitions): slice sampling with stepping out on a continuous site
        (infer/dispatch.py:190-194 -> infer/slice_sample.py:39-139)."""
        s, b = self._scope_block(scope, block)
        out = self._out()
        mv = self._out(dtype=np.int64) if return_moves else None
        _lib.call("vk_slice", self._h, s, b, float(w), float(m), int(transitions), _lib.ptr(out), _lib.ptr(mv))
        return (out, mv) if return_moves else out

    def slice_doubling(self, scope, block, w, p, transitions, return_moves=False):
        """(slice_doubling scope block w p transitions): slice sampling with doubling
        (infer/dispatch.py:195-199 -> DoublingSliceOperator, infer/slice_sample.py:148-196)."""
        s, b = self._scope_block(scope, block)
        out = self._out()
        mv = self._out(dtype=np.int64) if return_moves else None
        _lib.call("vk_slice_doubling", self._h, s, b, float(w), int(p), int(transitions), _lib.ptr(out), _lib.ptr(mv))
        return (out, mv) if return_moves else out

    def subsampled_mh(self, scope, block, Nbatch, k0, epsilon, useDeltaKernels, deltaKernelArgs,
                      updateValues, transitions):
        s, b = self._scope_block(scope, block)
        stats = np.empty((self.n_particles, 3))
        nodes = self._out()
        _lib.call("vk_subsampled_mh", self._h, s, b, int(Nbatch), int(k0), float(epsilon),
                  int(bool(useDeltaKernels)), float(deltaKernelArgs or 0.0), int(bool(updateValues)),
                  int(transitions), _lib.ptr(nodes), _lib.ptr(stats))
        self.last_mean_nodes = nodes       # what the reference's primitive_infer returns (dispatch.py:109-120, 149-152)
        return stats

    def log_likelihood_at(self, scope, block):
        s, b = self._scope_block(scope, block)
        out = self._out()
        _lib.call("vk_log_likelihood_at", self._h, s, b, _lib.ptr(out))
        return out

    def log_joint_at(self, scope, block):
        s, b = self._scope_block(scope, block)
        out = self._out()
        _lib.call("vk_log_joint_at", self._h, s, b, _lib.ptr(out))
        return out

    # -- state access --------------------------------------------------------
    def numBlocksInScope(self, scope="default"):
        n = ctypes.c_int64()
        _lib.call("vk_num_blocks", self._h, 0, ctypes.byref(n))
        return n.value

    def num_sites(self):
        n = ctypes.c_int64()
        _lib.call("vk_num_sites", self._h, ctypes.byref(n))
        return n.value

    def read_site(self, site):
        out = self._out()
        _lib.call("vk_read_site", self._h, int(site), _lib.ptr(out))
        return out

    def write_site(self, site, values):
        v = np.ascontiguousarray(values, dtype=np.float64)
        if v.shape != (self.n_particles,):
            raise VentureException("invalid_argument", "expected %d values" % self.n_particles)
        _lib.call("vk_write_site", self._h, int(site), _lib.ptr(v))

    def refresh(self):
        _lib.call("vk_refresh", self._h)

    def device_buffer(self, which):
        p, n = ctypes.c_void_p(), ctypes.c_int64()
        _lib.call("vk_device_buffer", self._h, int(which), ctypes.byref(p), ctypes.byref(n))
        return p.value, n.value
