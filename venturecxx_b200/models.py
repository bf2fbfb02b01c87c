"""Compiled fixed-structure models: what `constructScaffold`
(backend/lite/scaffold.py:149-173) would derive per transition is fixed at
model-compile time here; each class names the VentureScript program it stands
for and creates the matching device batch through the C ABI."""
import ctypes

import numpy as np

from . import _lib


class Model(object):
    kind = None

    def _create(self, n_particles, seed, device):
        raise NotImplementedError


class NormalNormalModel(Model):
    """assume x = normal(prior_mu, prior_sigma); observe normal(x, obs_sigma) = y_k
    (benchmarks/Makefile:6-9; test/inference_quality/micro/test_basic_stats.py:65-76)."""
    kind = 1

    def __init__(self, prior_mu=0.0, prior_sigma=1.0, obs_sigma=1.0, max_obs=1024):
        self.prior_mu, self.prior_sigma, self.obs_sigma = float(prior_mu), float(prior_sigma), float(obs_sigma)
        self.max_obs = int(max_obs)

    def _create(self, n_particles, seed, device):
        h = ctypes.c_void_p()
        _lib.call("vk_create_normal_normal", n_particles, seed, device, self.prior_mu, self.prior_sigma,
                  self.obs_sigma, self.max_obs, ctypes.byref(h))
        return h

    def _observe(self, h, y):
        y = np.ascontiguousarray(np.atleast_1d(y), dtype=np.float64)
        _lib.call("vk_nn_observe", h, _lib.ptr(y), y.size)


class LogisticRegressionModel(Model):
    """assume w_j = normal(prior_mu, prior_sigma) (j<D); assume w = array(w_0..);
    observe log_odds_flip(vector_dot(w, x_i)) = y_i   (SURVEY 8(d) C2/C5)."""
    kind = 2

    def __init__(self, D, n_obs_capacity, prior_mu=0.0, prior_sigma=1.0):
        self.D, self.n_obs_capacity = int(D), int(n_obs_capacity)
        self.prior_mu, self.prior_sigma = float(prior_mu), float(prior_sigma)

    def _create(self, n_particles, seed, device):
        h = ctypes.c_void_p()
        _lib.call("vk_create_blr", n_particles, seed, device, self.D, self.n_obs_capacity, self.prior_mu,
                  self.prior_sigma, ctypes.byref(h))
        return h

    def _observe(self, h, X, y):
        X = np.ascontiguousarray(X, dtype=np.float64).reshape(-1, self.D)
        y = np.ascontiguousarray(y).astype(np.uint8).reshape(-1)
        if X.shape[0] != y.shape[0]:
            raise _lib.VentureException("invalid_argument", "X and y disagree on the number of observations")
        _lib.call("vk_blr_observe", h, _lib.ptr(X), _lib.ptr(y), y.size)


class HMMModel(Model):
    """examples/hmm.vnt:31-60: state(1) = init_state, state(t) ~ categorical(T[state(t-1)-1], 1..S),
    observe categorical(E[state(t)-1], 1..O) = d_t."""
    kind = 3

    def __init__(self, transition_matrix, emission_matrix, init_state=1, T_max=32):
        self.trans = np.ascontiguousarray(transition_matrix, dtype=np.float64)
        self.emit = np.ascontiguousarray(emission_matrix, dtype=np.float64)
        self.S, self.O = self.emit.shape
        self.init_state, self.T_max = int(init_state), int(T_max)

    def _create(self, n_particles, seed, device):
        h = ctypes.c_void_p()
        _lib.call("vk_create_hmm", n_particles, seed, device, self.S, self.O, self.trans.shape[0],
                  _lib.ptr(self.trans), _lib.ptr(self.emit), self.init_state, self.T_max, ctypes.byref(h))
        return h

    def _observe(self, h, symbol):
        _lib.call("vk_hmm_observe", h, int(symbol))


class LDAModel(Model):
    """examples/lda.vnt:18-41 with collapsed make_sym_dir_cat makers and gamma hyperpriors."""
    kind = 4
    # scope "z": the topic draws alone, block = token index -- lda.vnt with ((doc doc_id)) wrapped in
    # (tag (quote z) <token> ...); needed by `gibbs` because the default scope also holds the gamma sites
    scopes = {"default": 0, "z": 1}

    def __init__(self, n_topics, vocab_size, n_docs, max_tokens, hyper_shape=1.0, hyper_rate=2.0):
        self.K, self.V, self.n_docs, self.max_tokens = int(n_topics), int(vocab_size), int(n_docs), int(max_tokens)
        self.hyper_shape, self.hyper_rate = float(hyper_shape), float(hyper_rate)

    def _create(self, n_particles, seed, device):
        h = ctypes.c_void_p()
        _lib.call("vk_create_lda", n_particles, seed, device, self.K, self.V, self.n_docs, self.max_tokens,
                  self.hyper_shape, self.hyper_rate, ctypes.byref(h))
        return h

    def _observe(self, h, doc, word):
        doc = np.ascontiguousarray(np.atleast_1d(doc), dtype=np.int32)
        word = np.ascontiguousarray(np.atleast_1d(word), dtype=np.int32)
        _lib.call("vk_lda_observe", h, _lib.ptr(doc), _lib.ptr(word), doc.size)
