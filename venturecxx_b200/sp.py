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


class CDirCatOutputPSP(B200PSP):
    """The SP made by make_dir_cat(alphas) (dirichlet.py:116-206); spaux() = counts (rows x n)."""
    sp_id = 13

    def _params(self, args):
        alpha = np.atleast_2d(np.asarray(args.operandValues()[0], dtype=np.float64))
        counts = np.atleast_2d(np.asarray(args.spaux(), dtype=np.float64))
        n = counts.shape[0]
        return np.ascontiguousarray(np.concatenate([np.broadcast_to(alpha, (n, counts.shape[1])), counts], axis=1)), n

    def _dim(self, params):
        return params.shape[1] // 2

    def logDensityOfData(self, args):
        params, n = self._params(args)
        out = np.empty(n)
        _lib.call("vk_sp_log_density_of_data", self.sp_id, _lib.ptr(params), params.shape[1], n, self._dim(params), args.device,
                  _lib.ptr(out))
        return out


class UDirCatOutputPSP(B200PSP):
    """The SP made by make_uc_dir_cat / make_uc_sym_dir_cat (dirichlet.py:265-293): operand = theta (rows x n)."""
    sp_id = 14

    def _params(self, args):
        theta = np.atleast_2d(np.asarray(args.operandValues()[0], dtype=np.float64))
        return np.ascontiguousarray(theta), theta.shape[0]

    def _dim(self, params):
        return params.shape[1]


class SuffBernoulliOutputPSP(B200PSP):
    """The SP made by make_uc_beta_bernoulli / make_suff_stat_bernoulli (discrete.py:513-545): operand = weight."""
    sp_id = 15

    def logDensityOfData(self, args):
        """spaux() = (yes, no) rows"""
        w = np.asarray(args.operandValues()[0], dtype=np.float64).reshape(-1, 1)
        aux = np.atleast_2d(np.asarray(args.spaux(), dtype=np.float64))
        n = aux.shape[0]
        params = np.ascontiguousarray(np.concatenate([np.broadcast_to(w, (n, 1)), aux], axis=1))
        out = np.empty(n)
        _lib.call("vk_sp_log_density_of_data", self.sp_id, _lib.ptr(params), 3, n, 1, args.device, _lib.ptr(out))
        return out


def _data_term(sp_id, params, dim, device):
    params = np.ascontiguousarray(params, dtype=np.float64)
    out = np.empty(params.shape[0])
    _lib.call("vk_sp_log_density_of_data", sp_id, _lib.ptr(params), params.shape[1], params.shape[0], dim, device, _lib.ptr(out))
    return out


class AAALKernel(object):
    """The AAA LKernels of the uncollapsed makers (lkernel.py:137-156 PosteriorAAALKernel): `simulate` draws the made
    SP's parameter from its conjugate posterior given the sufficient statistics, `weight` is the marginal
    likelihood of the data.  kind: "make_uc_dir_cat" (operands alphas; dirichlet.py:250-263), "make_uc_sym_dir_cat"
    (operands alpha, n; :395-407), "make_uc_beta_bernoulli" (operands a, b; discrete.py:502-510); spaux() = statistics."""
    KINDS = {"make_uc_dir_cat": 14, "make_uc_sym_dir_cat": 10, "make_uc_beta_bernoulli": 15}

    def __init__(self, kind):
        if kind not in self.KINDS:
            raise VentureException("invalid_argument", "%r has no AAA LKernel on the B200 path" % (kind,))
        self.kind, self.sp_id = kind, self.KINDS[kind]

    def _params(self, args):
        ops = [np.asarray(o, dtype=np.float64) for o in args.operandValues()]
        aux = np.atleast_2d(np.asarray(args.spaux(), dtype=np.float64))
        n = aux.shape[0]
        if self.kind == "make_uc_dir_cat":
            return np.concatenate([np.broadcast_to(np.atleast_2d(ops[0]), aux.shape), aux], axis=1), aux.shape[1], n
        if self.kind == "make_uc_sym_dir_cat":
            return np.concatenate([np.broadcast_to(ops[0].reshape(-1, 1), (n, 1)), aux], axis=1), aux.shape[1], n
        return np.concatenate([np.broadcast_to(ops[0].reshape(-1, 1), (n, 1)), np.broadcast_to(ops[1].reshape(-1, 1), (n, 1)), aux], axis=1), 1, n

    def simulate(self, args, return_weight=False):
        params, dim, n = self._params(args)
        params = np.ascontiguousarray(params)
        out = np.empty(n if self.kind == "make_uc_beta_bernoulli" else (n, dim))
        w = np.empty(n)
        _lib.call("vk_sp_aaa_simulate", self.sp_id, _lib.ptr(params), params.shape[1], n, dim, args.seed, args.step, args.device,
                  _lib.ptr(out), _lib.ptr(w))
        return (out, w) if return_weight else out

    def weight(self, args):
        params, dim, n = self._params(args)
        return _data_term({"make_uc_dir_cat": 13, "make_uc_sym_dir_cat": 11, "make_uc_beta_bernoulli": 12}[self.kind], params, dim, args.device)


builtins = {
    "normal": NormalOutputPSP, "gamma": GammaOutputPSP, "beta": BetaOutputPSP,
    "flip": BernoulliOutputPSP, "bernoulli": BernoulliOutputPSP, "log_flip": LogBernoulliOutputPSP,
    "log_bernoulli": LogBernoulliOutputPSP, "log_odds_flip": LogOddsBernoulliOutputPSP,
    "log_odds_bernoulli": LogOddsBernoulliOutputPSP, "categorical": CategoricalOutputPSP,
    "log_categorical": LogCategoricalOutputPSP, "dirichlet": DirichletOutputPSP,
    "symmetric_dirichlet": SymmetricDirichletOutputPSP, "make_sym_dir_cat": CSymDirCatOutputPSP,
    "make_beta_bernoulli": CBetaBernoulliOutputPSP, "make_dir_cat": CDirCatOutputPSP,
    "make_uc_dir_cat": UDirCatOutputPSP, "make_uc_sym_dir_cat": UDirCatOutputPSP,
    "make_uc_beta_bernoulli": SuffBernoulliOutputPSP, "make_suff_stat_bernoulli": SuffBernoulliOutputPSP,
}


def builtin_psp(name):
    try:
        return builtins[name]()
    except KeyError:
        raise VentureException("invalid_argument", "SP %r is not on the B200 path" % (name,))
