"""Shared test helpers: golden-fixture loading, oracle wrappers for the model
kernels (HMM, LDA), and the reference's statistical-test policy."""
import ctypes
import json
import math
import os

import numpy as np

import oracle

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def unj(x):
    if isinstance(x, list):
        return [unj(v) for v in x]
    if x == "nan":
        return float("nan")
    if x == "inf":
        return float("inf")
    if x == "-inf":
        return float("-inf")
    return x


def golden(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


def close(a, b, rtol=1e-12, atol=0.0):
    """equal incl. inf/nan, else |a-b| <= atol + rtol*|b|"""
    a, b = float(a), float(b)
    if math.isnan(a) or math.isnan(b):
        return math.isnan(a) and math.isnan(b)
    if math.isinf(a) or math.isinf(b):
        return a == b
    return abs(a - b) <= atol + rtol * abs(b)


# SP ids of include/venture_b200.h (enum vk_sp_id)
SP = dict(normal=1, gamma=2, beta=3, flip=4, log_flip=5, log_odds_flip=6, categorical=7, log_categorical=8,
          dirichlet=9, symmetric_dirichlet=10, sym_dir_cat=11, beta_bernoulli=12)


# ---- HMM oracle wrappers --------------------------------------------------
class OracleHMM(object):
    def __init__(self, trans, emit, init, P, Tpad, seed, p_base=0):
        self.trans = np.ascontiguousarray(trans, dtype=np.float64)
        self.emit = np.ascontiguousarray(emit, dtype=np.float64)
        self.S, self.O = self.emit.shape
        self.init, self.P, self.Tpad, self.seed, self.p_base = init, P, Tpad, seed, p_base
        self.hist = np.zeros((P, Tpad), dtype=np.uint8)
        self.lw = np.zeros(P)
        self.T = 0
        self.symbols = []

    def step(self, symbol, rng_step):
        incr = np.empty(self.P)
        oracle.lib().orc_hmm_step(oracle._ptr(self.hist), oracle.I32(self.Tpad), oracle.I32(self.T), oracle._ptr(self.lw),
                                  oracle._ptr(incr), oracle._ptr(self.trans), oracle._ptr(self.emit), oracle.I32(self.S),
                                  oracle.I32(self.O), oracle.I32(self.init), oracle.I32(symbol), oracle.I64(self.P),
                                  oracle.I64(self.p_base), oracle.U64(self.seed), oracle.U32(rng_step))
        self.T += 1
        self.symbols.append(symbol)
        return incr

    def prior(self, rng_step, write_lw=True):
        sym = np.ascontiguousarray(self.symbols, dtype=np.int32)
        oracle.lib().orc_hmm_prior(oracle._ptr(self.hist), oracle.I32(self.Tpad), oracle.I32(self.T),
                                   oracle._ptr(self.lw) if write_lw else None, oracle._ptr(self.trans),
                                   oracle._ptr(self.emit), oracle.I32(self.S), oracle.I32(self.O), oracle.I32(self.init),
                                   oracle._ptr(sym), oracle.I64(self.P), oracle.I64(self.p_base), oracle.U64(self.seed),
                                   oracle.U32(rng_step))

    def mh(self, step0, n_steps, block_mode):
        sym = np.ascontiguousarray(self.symbols, dtype=np.int32)
        acc = np.empty(self.P, dtype=np.int64)
        oracle.lib().orc_hmm_mh(oracle._ptr(self.hist), oracle.I32(self.Tpad), oracle.I32(self.T), oracle._ptr(self.trans),
                                oracle._ptr(self.emit), oracle.I32(self.S), oracle.I32(self.O), oracle._ptr(sym),
                                oracle.I64(self.P), oracle.I64(self.p_base), oracle.U64(self.seed), oracle.U32(step0),
                                oracle.I32(n_steps), oracle.I32(block_mode), oracle._ptr(acc))
        return acc

    def gibbs(self, step0, n_steps, block_mode):
        sym = np.ascontiguousarray(self.symbols, dtype=np.int32)
        ch = np.empty(self.P, dtype=np.int64)
        oracle.lib().orc_hmm_gibbs(oracle._ptr(self.hist), oracle.I32(self.Tpad), oracle.I32(self.T), oracle._ptr(self.trans),
                                   oracle._ptr(self.emit), oracle.I32(self.S), oracle.I32(self.O), oracle._ptr(sym),
                                   oracle.I64(self.P), oracle.I64(self.p_base), oracle.U64(self.seed), oracle.U32(step0),
                                   oracle.I32(n_steps), oracle.I32(block_mode), oracle._ptr(ch))
        return ch

    def pgibbs(self, step0, n_steps, particles):
        sym = np.ascontiguousarray(self.symbols, dtype=np.int32)
        acc = np.empty(self.P, dtype=np.int64)
        oracle.lib().orc_hmm_pgibbs(oracle._ptr(self.hist), oracle.I32(self.Tpad), oracle.I32(self.T), oracle._ptr(self.trans),
                                    oracle._ptr(self.emit), oracle.I32(self.S), oracle.I32(self.O), oracle._ptr(sym),
                                    oracle.I64(self.P), oracle.I64(self.p_base), oracle.U64(self.seed), oracle.U32(step0),
                                    oracle.I32(n_steps), oracle.I32(particles), oracle._ptr(acc))
        return acc

    def resample(self, anc):
        self.hist = np.ascontiguousarray(self.hist[anc])
        lse = oracle.logsumexp(self.lw)
        self.lw = np.full(self.P, lse - math.log(self.P))


def hmm_forward_marginals(trans, emit, data, init):
    """Exact forward filter (numpy) -- the quantity examples/hmm_plugin.py:175-205 tabulates."""
    trans, emit = np.asarray(trans)[:len(emit)], np.asarray(emit)
    trans = trans / trans.sum(1, keepdims=True)
    emit = emit / emit.sum(1, keepdims=True)
    S = emit.shape[0]
    out = []
    a = np.zeros(S); a[init - 1] = 1.0
    for t, d in enumerate(data):
        if t > 0:
            a = a @ trans
        a = a * emit[:, d - 1]
        a = a / a.sum()
        out.append(a.copy())
    return np.array(out)


# ---- LDA oracle wrappers --------------------------------------------------
class _OrcLDA(ctypes.Structure):
    _fields_ = [("K", ctypes.c_int), ("V", ctypes.c_int), ("n_docs", ctypes.c_int), ("max_tok", ctypes.c_int),
                ("a", ctypes.c_double), ("b", ctypes.c_double)]


class OracleLDA(object):
    def __init__(self, K, V, n_docs, max_tok, a, b, P, seed, p_base=0):
        self.L = _OrcLDA(K, V, n_docs, max_tok, a, b)
        self.K, self.V, self.n_docs, self.max_tok, self.P, self.seed, self.p_base = K, V, n_docs, max_tok, P, seed, p_base
        self.alpha = np.zeros((P, 2))
        self.nk = np.zeros((P, K), dtype=np.int32)
        self.ndk = np.zeros((P, n_docs * K), dtype=np.int32)
        self.nkw = np.zeros((P, K * V), dtype=np.int32)
        self.z = np.zeros((P, max_tok), dtype=np.uint8)
        self.lw = np.zeros(P)
        self.doc = np.zeros(max_tok, dtype=np.int32)
        self.word = np.zeros(max_tok, dtype=np.int32)
        self.n_tok = 0

    def observe(self, doc, word):
        n = len(doc)
        self.doc[self.n_tok:self.n_tok + n] = doc
        self.word[self.n_tok:self.n_tok + n] = word
        self.n_tok += n

    def seq(self, mode, t0, t1, rng_step):
        incr = np.zeros(self.P)
        oracle.lib().orc_lda_seq(oracle._ptr(self.alpha), oracle._ptr(self.nk), oracle._ptr(self.ndk), oracle._ptr(self.nkw),
                                 oracle._ptr(self.z), ctypes.byref(self.L), oracle._ptr(self.doc), oracle._ptr(self.word),
                                 oracle.I32(t0), oracle.I32(t1), oracle._ptr(self.lw), oracle._ptr(incr), oracle.I64(self.P),
                                 oracle.I64(self.p_base), oracle.U64(self.seed), oracle.U32(rng_step), oracle.I32(mode))
        return incr

    def mh(self, step0, n_steps, block_mode):
        acc = np.empty(self.P, dtype=np.int64)
        oracle.lib().orc_lda_mh(oracle._ptr(self.alpha), oracle._ptr(self.nk), oracle._ptr(self.ndk), oracle._ptr(self.nkw),
                                oracle._ptr(self.z), ctypes.byref(self.L), oracle._ptr(self.doc), oracle._ptr(self.word),
                                oracle.I32(self.n_tok), oracle.I64(self.P), oracle.I64(self.p_base), oracle.U64(self.seed),
                                oracle.U32(step0), oracle.I32(n_steps), oracle.I32(block_mode), oracle._ptr(acc))
        return acc

    def gibbs(self, step0, n_steps, block_mode):
        """Enumerative Gibbs over the z sites (blocks = tokens)."""
        ch = np.empty(self.P, dtype=np.int64)
        oracle.lib().orc_lda_gibbs(oracle._ptr(self.alpha), oracle._ptr(self.nk), oracle._ptr(self.ndk), oracle._ptr(self.nkw),
                                   oracle._ptr(self.z), ctypes.byref(self.L), oracle._ptr(self.doc), oracle._ptr(self.word),
                                   oracle.I32(self.n_tok), oracle.I64(self.P), oracle.I64(self.p_base), oracle.U64(self.seed),
                                   oracle.U32(step0), oracle.I32(n_steps), oracle.I32(block_mode), oracle._ptr(ch))
        return ch
