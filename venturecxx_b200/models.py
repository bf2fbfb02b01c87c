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

    # -- interchange.py: the program this model stands for, and its random values in evaluation order
    def _ys(self, obs):
        return [float(v) for (y,) in obs for v in np.atleast_1d(y)]

    def program(self, obs):
        from .interchange import define, number, observe
        return [define("x", ["normal", self.prior_mu, self.prior_sigma])] + \
            [observe(["normal", "x", self.obs_sigma], number(y)) for y in self._ys(obs)]

    def random_values(self, obs, sites):
        from .interchange import number
        return [number(sites[0])] + [number(y) for y in self._ys(obs)]

    def sites_from_random_values(self, obs, values):
        from .interchange import unbox
        return np.array([unbox(values[0])])


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

    # subsampled MH (backend/lite/infer/subsampled_mh.py:326-368): every weight's global border is the `(array w_0 ...)`
    # node and its children are the N `vector_dot` applications, one local section per observation
    def subsampled_applicability(self, block):
        return True, None, ""

    def subsampled_global_nodes(self, block):
        return 3.0 * (self.D if block in ("one", "each", -1, -3) else 1)      # per block: the weight, its lookup, the array

    def _rows(self, obs):
        for X, y in obs:
            X = np.asarray(X, dtype=np.float64).reshape(-1, self.D)
            y = np.asarray(y).reshape(-1)
            for i in range(len(y)):
                yield X[i], bool(y[i])

    def program(self, obs):
        from .interchange import boolean, define, observe
        prog = [define("w_%d" % j, ["normal", self.prior_mu, self.prior_sigma]) for j in range(self.D)]
        prog.append(define("w", ["array"] + ["w_%d" % j for j in range(self.D)]))
        for x, y in self._rows(obs):
            prog.append(observe(["log_odds_flip", ["vector_dot", "w", ["array"] + [float(v) for v in x]]], boolean(y)))
        return prog

    def random_values(self, obs, sites):
        from .interchange import boolean, number
        return [number(sites[j]) for j in range(self.D)] + [boolean(y) for _x, y in self._rows(obs)]

    def sites_from_random_values(self, obs, values):
        from .interchange import unbox
        return np.array([unbox(values[j]) for j in range(self.D)])


class HMMModel(Model):
    """examples/hmm.vnt:31-60: state(1) = init_state, state(t) ~ categorical(T[state(t-1)-1], 1..S),
    observe categorical(E[state(t)-1], 1..O) = d_t."""
    kind = 3
    # hmm.vnt tags every state(t) into the ordered scope `seq`, block t: the same random choices as the default scope
    scopes = {"default": 0, "seq": 0}

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

    def program(self, obs):
        """examples/hmm.vnt:31-60 and the observe of :71; (state t) requests state(t-1) recursively
        through mem, so observing times 1..T in order creates the states in order."""
        from .interchange import define, number, observe
        states = ["array"] + list(range(1, self.S + 1))
        outs = ["array"] + list(range(1, self.O + 1))
        prog = [define("transition_matrix", ["array"] + [["simplex"] + [float(v) for v in r] for r in self.trans]),
                define("emission_matrix", ["array"] + [["simplex"] + [float(v) for v in r] for r in self.emit]),
                define("transition", ["lambda", ["state"],
                                      ["categorical", ["lookup", "transition_matrix", ["sub", "state", 1]], states]]),
                define("emission", ["lambda", ["state"],
                                    ["categorical", ["lookup", "emission_matrix", ["sub", "state", 1]], outs]]),
                define("state", ["mem", ["lambda", ["t"], ["tag", ["quote", "seq"], "t",
                                                           ["if", ["lte", "t", 1], self.init_state,
                                                            ["transition", ["state", ["sub", "t", 1]]]]]]])]
        for t, (d,) in enumerate(obs):
            prog.append(observe(["emission", ["state", t + 1]], number(int(d))))
        return prog

    def random_values(self, obs, sites):
        # evaluation order: obs_1 (state(1) is deterministic); then state(t), obs_t for t = 2..T
        from .interchange import number
        out = []
        for t, (d,) in enumerate(obs):
            if t > 0:
                out.append(number(int(sites[t])))
            out.append(number(int(d)))
        return out

    def sites_from_random_values(self, obs, values):
        from .interchange import unbox
        T = len(obs)
        sites = np.empty(T)
        if T:
            sites[0] = self.init_state
        for t in range(1, T):
            sites[t] = unbox(values[2 * t - 1])
        return sites


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

    def _tokens(self, obs):
        pos = {}
        for doc, word in obs:
            for d, w in zip(np.atleast_1d(doc), np.atleast_1d(word)):
                d = int(d)
                pos[d] = pos.get(d, -1) + 1
                yield d, pos[d], int(w)

    def program(self, obs):
        """examples/lda.vnt:18-41: (model n_topics vocab_size) then one observe per token."""
        from .interchange import define, integer, observe
        prog = [define("n_topics", self.K), define("vocab_size", self.V),
                define("alpha_document_topic", ["gamma", self.hyper_shape, self.hyper_rate]),
                define("alpha_topic_word", ["gamma", self.hyper_shape, self.hyper_rate]),
                define("doc", ["mem", ["lambda", ["doc_id"], ["make_sym_dir_cat", "alpha_document_topic", "n_topics"]]]),
                define("topic", ["mem", ["lambda", ["topic_id"], ["make_sym_dir_cat", "alpha_topic_word", "vocab_size"]]]),
                define("word", ["mem", ["lambda", ["doc_id", "pos"], [["topic", [["doc", "doc_id"]]]]]])]
        for d, pos, w in self._tokens(obs):
            prog.append(observe(["word", d, pos], integer(w)))
        return prog

    def random_values(self, obs, sites):
        # evaluation order: the two gamma assumes; per token the topic draw ((doc d)) -- a VentureInteger,
        # the default outcome objects of make_sym_dir_cat (dirichlet.py:218) -- then the observed word
        from .interchange import integer, number
        out = [number(sites[0]), number(sites[1])]
        for i, (_d, _pos, w) in enumerate(self._tokens(obs)):
            out.append(integer(sites[2 + i]))
            out.append(integer(w))
        return out

    def sites_from_random_values(self, obs, values):
        from .interchange import unbox
        n = sum(1 for _ in self._tokens(obs))
        return np.array([unbox(values[0]), unbox(values[1])] + [unbox(values[2 + 2 * i]) for i in range(n)], dtype=np.float64)
