"""CPU oracle for the particle inference path -- TEST INFRASTRUCTURE ONLY.

Importable from tests/, __graft_entry__.smoke() and bench.py's CPU-baseline
legs.  The product package (venturecxx_b200) never imports this module.

  oracle.c            plain-C restatement consuming the product's Philox
                      streams (exact-chain comparisons, timed CPU baseline)
  lite_semantics.py   numpy/scipy restatement of the Lite functions, pinned
                      against the reference's golden vectors
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

c_double_p = ctypes.POINTER(ctypes.c_double)


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle.so"],
                              stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(so):
            build()
        L = ctypes.CDLL(so)
        L.orc_logsumexp.restype = ctypes.c_double
        L.orc_log_logistic.restype = ctypes.c_double
        L.orc_logistic.restype = ctypes.c_double
        L.orc_log_logistic.argtypes = [ctypes.c_double]
        L.orc_logistic.argtypes = [ctypes.c_double]
        L.orc_sym_dircat_log_density_of_data.restype = ctypes.c_double
        L.orc_beta_bernoulli_log_density_of_data.restype = ctypes.c_double
        L.orc_beta_bernoulli_log_density_of_data.argtypes = [ctypes.c_double] * 4
        L.orc_student_t_cdf.restype = ctypes.c_double
        L.orc_student_t_cdf.argtypes = [ctypes.c_double, ctypes.c_double]
        _LIB = L
    return _LIB


def _d(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a, a.ctypes.data_as(c_double_p)


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


I64 = ctypes.c_int64
U64 = ctypes.c_uint64
U32 = ctypes.c_uint32
I32 = ctypes.c_int32
F64 = ctypes.c_double


def uniforms(seed, particle, step, n):
    out = np.empty(n)
    lib().orc_uniforms(U64(seed), U64(particle), U32(step), I32(n), _ptr(out))
    return out


def logsumexp(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return lib().orc_logsumexp(_ptr(a), I64(a.size))


def normalized_probs(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    out = np.empty_like(a)
    lib().orc_normalized_probs(_ptr(a), I64(a.size), _ptr(out))
    return out


def ancestors(lw, n_new, mode, seed=0, step=0, uniforms=None):
    lw = np.ascontiguousarray(lw, dtype=np.float64)
    anc = np.empty(n_new, dtype=np.int32)
    u = None if uniforms is None else np.ascontiguousarray(uniforms, dtype=np.float64)
    lib().orc_ancestors(_ptr(lw), I64(lw.size), I64(n_new), I32(mode), U64(seed), U32(step),
                        _ptr(u), _ptr(anc))
    return anc


def sp_simulate(sp, params, n, dim=1, seed=0, step=0):
    params = np.ascontiguousarray(params, dtype=np.float64).reshape(n, -1)
    out = np.empty((n, dim) if dim > 1 and sp in (9, 10) else n)
    rc = lib().orc_sp_simulate(I32(sp), _ptr(params), I32(params.shape[1]), I64(n), I32(dim),
                               U64(seed), U32(step), _ptr(out))
    assert rc == 0
    return out


def sp_log_density(sp, values, params, n, dim=1):
    params = np.ascontiguousarray(params, dtype=np.float64).reshape(n, -1)
    values = np.ascontiguousarray(values, dtype=np.float64)
    out = np.empty(n)
    rc = lib().orc_sp_log_density(I32(sp), _ptr(values), _ptr(params), I32(params.shape[1]), I64(n),
                                  I32(dim), _ptr(out))
    assert rc == 0
    return out
