"""Batched mirror of the reference's PSP interface (backend/lite/psp.py:32-318)
for the builtins on the hot path.  `simulate(args)` / `logDensity(value, args)`
keep their names and meaning; `args.operandValues()` returns one ARRAY per
operand (one application per row) instead of one value, and the work runs in
the CUDA library (vk_sp_simulate / vk_sp_log_density).  The registry is keyed
by the names `registerBuiltinSP` uses in continuous.py / discrete.py /
dirichlet.py."""
import numpy as np

from . import _lib
from .exception import VentureException


class BatchArgs(object):
    """IArgs (psp.py:320-383) for a batch: operandValues() plus the Philox address
    (seed, step) that replaces np_prng()/py_prng(); spaux() carries count
    statistics for the collapsed SPs."""

    def __init__(self, operands, seed=0, step=0, aux=None, device=0):
        self._operands = [np.atleast_1d(np.asarray(o, dtype=np.float64)) for o in operands]
        self.seed, self.step, self._aux, self.device = int(seed), int(step), aux, int(device)

    def operandValues(self):
        return self._operands

    def spaux(self):
        return self._aux


class B200PSP(object):
    sp_id = None
    vector_valued = False

    def isRandom(self):
        return True

    def canAbsorb(self, _trace=None, _appNode=None, _parentNode=None):
        return True  # RandomPSP.canAbsorb, psp.py:459-461

    def _params(self, args):
        ops = args.operandValues()
        n = max(o.shape[0] for o in ops)
        cols = [np.broadcast_to(o.reshape(o.shape[0], -1), (n, int(np.prod(o.shape[1:])) or 1)) for o in ops]
        return np.ascontiguousarray(np.concatenate(cols, axis=1)), n

    def _dim(self, params):
        return 1

    def simulate(self, args):
        params, n = self._params(args)
        dim = self._dim(params)
        out = np.empty((n, dim) if self.vector_valued else n)
        _lib.call("vk_sp_simulate", self.sp_id, _lib.ptr(params), params.shape[1], n, dim, args.seed, args.step,
                  args.device, _lib.ptr(out))
        return out

    def logDensity(self, value, args):
        params, n = self._params(args)
        dim = self._dim(params)
        v = np.ascontiguousarray(np.broadcast_to(np.asarray(value, dtype=np.float64),
                                                 (n, dim) if self.vector_valued else (n,)))
        out = np.empty(n)
        _lib.call("vk_sp_log_density", self.sp_id, _lib.ptr(v), _lib.ptr(params), params.shape[1], n, dim, args.device,
                  _lib.ptr(out))
        return out


def _psp(name, sp_id, vector=False, dim_fn=None):
    cls = type(name, (B200PSP,), {"sp_id": sp_id, "vector_valued": vector})
    if dim_fn:
        cls._dim = lambda self, params: dim_fn(params)
    return cls


NormalOutputPSP = _psp("NormalOutputPSP", 1)                # continuous.py:313-375
GammaOutputPSP = _psp("GammaOutputPSP", 2)                  # continuous.py:1161-1226
BetaOutputPSP = _psp("BetaOutputPSP", 3)                    # continuous.py:729-882
BernoulliOutputPSP = _psp("BernoulliOutputPSP", 4)          # discrete.py:55-102
LogBernoulliOutputPSP = _psp("LogBernoulliOutputPSP", 5)    # discrete.py:105-140
LogOddsBernoulliOutputPSP = _psp("LogOddsBernoulliOutputPSP", 6)  # discrete.py:143-178
CategoricalOutputPSP = _psp("CategoricalOutputPSP", 7)      # discrete.py:208-246 (default objects 0..n-1)
LogCategoricalOutputPSP = _psp("LogCategoricalOutputPSP", 8)  # discrete.py:249-287
DirichletOutputPSP = _psp("DirichletOutputPSP", 9, True, lambda p: p.shape[1])  # dirichlet.py:41-57


class SymmetricDirichletOutputPSP(B200PSP):                 # dirichlet.py:59-77: operands (alpha, n)
    sp_id, vector_valued = 10, True

    def _params(self, args):
        alpha, n_dim = args.operandValues()
        self._n = int(n_dim[0])
        return np.ascontiguousarray(alpha.reshape(-1, 1)), alpha.shape[0]

    def _dim(self, params):
        return self._n


class CSymDirCatOutputPSP(B200PSP):
    """The SP made by make_sym_dir_cat(alpha, n) (dirichlet.py:158-201, 303-351); spaux() = counts (rows x n)."""
    sp_id = 11

    def _params(self, args):
        alpha = args.operandValues()[0]
        counts = np.atleast_2d(np.asarray(args.spaux(), dtype=np.float64))
        n = counts.shape[0]
        return np.ascontiguousarray(np.concatenate([np.broadcast_to(alpha.reshape(-1, 1), (n, 1)), counts], axis=1)), n

    def _dim(self, params):
        return params.shape[1] - 1


class CBetaBernoulliOutputPSP(B200PSP):
    """The SP made by make_beta_bernoulli(a, b) (discrete.py:392-463); spaux() = (yes, no) rows."""
    sp_id = 12

    def _params(self, args):
        a, b = args.operandValues()
        aux = np.atleast_2d(np.asarray(args.spaux(), dtype=np.float64))
        n = aux.shape[0]
        return np.ascontiguousarray(np.concatenate([np.broadcast_to(a.reshape(-1, 1), (n, 1)),
                                                    np.broadcast_to(b.reshape(-1, 1), (n, 1)), aux], axis=1)), n


builtins = {
    "normal": NormalOutputPSP, "gamma": GammaOutputPSP, "beta": BetaOutputPSP,
    "flip": BernoulliOutputPSP, "bernoulli": BernoulliOutputPSP, "log_flip": LogBernoulliOutputPSP,
    "log_bernoulli": LogBernoulliOutputPSP, "log_odds_flip": LogOddsBernoulliOutputPSP,
    "log_odds_bernoulli": LogOddsBernoulliOutputPSP, "categorical": CategoricalOutputPSP,
    "log_categorical": LogCategoricalOutputPSP, "dirichlet": DirichletOutputPSP,
    "symmetric_dirichlet": SymmetricDirichletOutputPSP, "make_sym_dir_cat": CSymDirCatOutputPSP,
    "make_beta_bernoulli": CBetaBernoulliOutputPSP,
}


def builtin_psp(name):
    try:
        return builtins[name]()
    except KeyError:
        raise VentureException("invalid_argument", "SP %r is not on the B200 path" % (name,))
