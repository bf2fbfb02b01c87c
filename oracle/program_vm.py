"""TEST INFRASTRUCTURE: a pure-Python interpreter of the model compiler's register-machine programs
(venturecxx_b200/compiler.py), for small cases.  It executes the SAME blob the device gets, with

  * draws from the oracle's samplers at the matching positions of the Philox streams (oracle.draw -> orc_draw,
    the same stand-ins tests/golden/make_program_chains.py puts behind Lite's np_rng), and
  * log densities from the reference's own call sites: normal by Lite's formula (continuous.py:319-323),
    gamma / beta / uniform / expon through scipy.stats as Lite does (continuous.py:1166-1167, 858-859, 639, 1137),
    the flip family by discrete.py:61-67, 109-113, 148-150 with utils.py:60-61, 257-268,

so that tests can pin the COMPILER (which nodes absorb where, in which order values are simulated, the
transition protocol of infer/mh.py:59-113, 167-208) against the Lite fixture without a GPU.
"""
import math

import numpy as np
import scipy.stats

import oracle

WORDS = 4


def _log(x):
    return -math.inf if x == 0 else math.log(x)


def _log1p(x):
    return -math.inf if x == -1 else math.log1p(x)


def _log_logistic(x):
    return x if x <= -37 else -math.log1p(math.exp(-x))


def _logistic(x):
    return math.exp(x) if x <= -37 else 1.0 / (1.0 + math.exp(-x))


def _normalize(seq):                                          # normalizeList, utils.py:93-99
    den = sum(seq)
    if den > 0:
        return [float(x) / den for x in seq]
    return [1.0 / len(seq) for _ in seq]


class VM(object):
    def __init__(self, blob, op_table, P, seed, p_base=0):
        self.code, self.consts, meta, self.data, self.affected = [np.asarray(b) for b in blob]
        self.op = dict((v, k) for k, v in op_table.items())
        self.P, self.seed, self.p_base = P, seed, p_base
        m = [int(x) for x in meta]
        self.S, self.G, self.n_regs, self.n_sec, self.sec_prior, self.sec_joint = m[:6]
        at = 6
        self.site = [m[at + 4 * k: at + 4 * k + 4] for k in range(self.S)]
        at += 4 * self.S
        self.site_groups = []
        for k in range(self.S):
            self.site_groups.append(m[at: at + self.site[k][3]])
            at += self.site[k][3]
        self.group = [m[at + 5 * g: at + 5 * g + 5] for g in range(self.G)]
        at += 5 * self.G
        self.dir = [(m[at + 2 * s], m[at + 2 * s + 1]) for s in range(self.n_sec)]
        at += 2 * self.n_sec
        self.ss = [m[at + 3 * k: at + 3 * k + 3] for k in range(self.S)]          # kind, N, offset into the roots
        at += 3 * self.S
        n_roots = m[at]
        at += 1
        self.roots = [(m[at + 2 * i], m[at + 2 * i + 1]) for i in range(n_roots)]
        at += 2 * n_roots
        n_scopes = m[at]
        at += 1
        tab = [(m[at + 2 * i], m[at + 2 * i + 1]) for i in range(n_scopes)]
        at += 2 * n_scopes
        self.scope_sites = [m[at + off: at + off + n] for n, off in tab]       # scope 0 = default: every site
        at += sum(n for n, _ in tab)
        ctab = [(m[at + 2 * k], m[at + 2 * k + 1]) for k in range(self.S)]
        at += 2 * self.S
        self.candidates = [[float(self.consts[m[at + off + j]]) for j in range(n)] for n, off in ctab]
        self.sites = np.zeros((self.S, P))
        self.lw = np.zeros(P)
        self.step = 0

    # -- one section ------------------------------------------------------------------------------------------
    def run(self, sec, R, p, row=None, tape=None, store=True):
        first, n = self.dir[sec]
        acc = 0.0
        for i in range(first, first + n):
            op, d, a, b = [int(x) for x in self.code[i * WORDS:(i + 1) * WORDS]]
            name = self.op[op]
            if name == "CONST": R[d] = float(self.consts[a])
            elif name == "SITE": R[d] = float(self.sites[a, p])
            elif name == "DATA": R[d] = float(row[a])
            elif name == "ADD": R[d] = R[a] + R[b]
            elif name == "SUB": R[d] = R[a] - R[b]
            elif name == "MUL": R[d] = R[a] * R[b]
            elif name == "DIV": R[d] = R[a] / R[b]
            elif name == "NEG": R[d] = -R[a]
            elif name == "EXP": R[d] = math.exp(R[a])
            elif name == "LOG": R[d] = _log(R[a])
            elif name == "SQRT": R[d] = math.sqrt(R[a])
            elif name == "POW": R[d] = math.pow(R[a], R[b])
            elif name == "LOGISTIC": R[d] = _logistic(R[a])
            elif name == "ABS": R[d] = abs(R[a])
            elif name == "MIN": R[d] = min(R[a], R[b])
            elif name == "MAX": R[d] = max(R[a], R[b])
            elif name == "LOG1P": R[d] = _log1p(R[a])
            elif name == "EXPM1": R[d] = math.expm1(R[a])
            elif name == "LOG_LOGISTIC": R[d] = _log_logistic(R[a])
            elif name == "LOGIT": R[d] = math.log(R[a] / (1 - R[a]))
            elif name == "MOV": R[d] = R[a]
            elif name == "ARGN": self._argn = a
            elif name == "LOOKUP":
                j = int(R[b])
                R[d] = R[a + j] if 0 <= j < self._argn else math.nan
            elif name == "SIM_CATEGORICAL":                   # simulateCategorical, utils.py:101-104: normalise, one draw
                k = self._argn
                ps = _normalize([R[a + j] for j in range(k)])
                R[d] = R[a + k + int(tape(3, ps))]
            elif name == "LD_CATEGORICAL":                    # logDensityCategorical, utils.py:106-118
                k = self._argn
                ps = _normalize([R[a + j] for j in range(k)])
                pm = sum(pj for j, pj in enumerate(ps) if R[a + k + j] == R[d])
                acc += -math.inf if pm == 0 else math.log(pm)
            elif name == "COND": self._flag = R[a] != 0
            elif name == "SELECT": R[d] = R[a] if self._flag else R[b]
            elif name == "SIM_NORMAL": R[d] = tape(1, [R[a], R[b]])
            elif name == "SIM_GAMMA": R[d] = tape(2, [R[a], R[b]])
            elif name == "SIM_BETA": R[d] = tape(5, [R[a], R[b]])
            elif name == "SIM_UNIFORM": R[d] = R[a] + (R[b] - R[a]) * tape(0, [0.0])
            elif name == "SIM_EXPON": R[d] = -math.log1p(-tape(0, [0.0])) * (1.0 / R[a])
            elif name == "SIM_FLIP": R[d] = 1.0 if tape(0, [0.0]) < R[a] else 0.0
            elif name == "SIM_LOG_FLIP": R[d] = 1.0 if math.log(tape(0, [0.0])) < R[a] else 0.0
            elif name == "SIM_LOG_ODDS_FLIP":
                u = tape(0, [0.0])
                R[d] = 1.0 if math.log(u / (1 - u)) < R[a] else 0.0
            elif name == "LD_NORMAL":
                dev = R[d] - R[a]                       # continuous.py:319-323
                acc += -math.log(R[b]) - 0.5 * math.log(2 * math.pi) - 0.5 * dev * dev / (R[b] * R[b])
            elif name == "LD_GAMMA": acc += float(scipy.stats.gamma.logpdf(R[d], R[a], scale=1.0 / R[b]))
            elif name == "LD_BETA": acc += float(scipy.stats.beta.logpdf(R[d], R[a], R[b]))
            elif name == "LD_UNIFORM": acc += float(scipy.stats.uniform.logpdf(R[d], R[a], R[b] - R[a]))
            elif name == "LD_EXPON": acc += float(scipy.stats.expon.logpdf(R[d], scale=1.0 / R[a]))
            elif name == "LD_FLIP": acc += _log(R[a]) if R[d] != 0 else _log1p(-R[a])
            elif name == "LD_LOG_FLIP": acc += R[a] if R[d] != 0 else _log1p(-math.exp(R[a]))
            elif name == "LD_LOG_ODDS_FLIP": acc += _log_logistic(R[a] if R[d] != 0 else -R[a])
            elif name == "STORE":
                if store:
                    self.sites[a, p] = R[d]
            else:
                raise ValueError("unknown op %d" % op)
        return acc

    def _tape(self, p, step, d0):
        state = {"d": d0}

        def draw(kind, params):
            v, state["d"] = oracle.draw(kind, params, self.seed, self.p_base + p, step, state["d"])
            return v
        return draw

    def group_ll(self, g, R, p):
        pro, row_sec, n_rows, n_cols, off = self.group[g]
        self.run(pro, R, p)
        tot = 0.0
        for r in range(n_rows):
            tot += self.run(row_sec, R, p, row=self.data[off + r * n_cols: off + (r + 1) * n_cols])
        return tot

    def children(self, k, R, p):
        acc = self.run(self.site[k][1], R, p)
        for g in self.site_groups[k]:
            acc += self.group_ll(g, R, p)
        return acc

    def all_groups(self, R, p):
        return sum(self.group_ll(g, R, p) for g in range(self.G))

    def pick(self, u_sel, t, block, scope):
        """block -1: one (uniform over the scope's blocks), -3: each (in order), k >= 0: block k of the scope -> site"""
        sites = self.scope_sites[scope]
        if block == -1:
            return sites[min(int(u_sel * len(sites)), len(sites) - 1)]
        if block == -3:
            return sites[t % len(sites)]
        return sites[block]

    # -- the operations of the C ABI --------------------------------------------------------------------------
    def likelihood_weight(self):
        step, self.step = self.step, self.step + 1
        for p in range(self.P):
            R = [0.0] * self.n_regs
            self.run(self.sec_prior, R, p, tape=self._tape(p, step, 0))
            self.lw[p] = self.all_groups(R, p)

    def reset_to_prior(self):
        step, self.step = self.step, self.step + 1
        for p in range(self.P):
            self.run(self.sec_prior, [0.0] * self.n_regs, p, tape=self._tape(p, step, 0))

    def log_likelihood(self, joint=False):
        out = np.zeros(self.P)
        for p in range(self.P):
            R = [0.0] * self.n_regs
            out[p] = self.all_groups(R, p) + (self.run(self.sec_joint, R, p) if joint else 0.0)
        return out

    def mh(self, n_steps, block=-1, scope=0):
        """block -1: one, -3: each, k >= 0: block k of the scope.  Returns (accept counts, per-particle mean node counts)."""
        step0, self.step = self.step, self.step + n_steps
        acc, nodes = np.zeros(self.P, dtype=np.int64), np.zeros(self.P)
        for p in range(self.P):
            for t in range(n_steps):
                u = oracle.uniforms(self.seed, self.p_base + p, step0 + t, 2)
                k = self.pick(u[0], t, block, scope)
                R = [0.0] * self.n_regs
                rho = self.children(k, R, p)
                old = self.sites[k, p]
                self.run(self.site[k][0], R, p, tape=self._tape(p, step0 + t, 2))
                xi = self.children(k, R, p)
                if rho == -math.inf and xi == -math.inf:
                    alpha = math.inf
                elif rho == math.inf and xi == math.inf:
                    alpha = -math.inf
                else:
                    alpha = xi - rho
                if math.log(u[1]) < alpha:
                    acc[p] += 1
                else:
                    self.sites[k, p] = old
                nodes[p] += self.affected[k]
        return acc, nodes / max(n_steps, 1)

    def gibbs(self, n_steps, block=-1, scope=0, map=False):
        """EnumerativeGibbsOperator under mixMH (infer/egibbs.py:42-107) on a boolean site: candidates [True, False]
        (discrete.py:78-86), each weighted by the regen weight with the value forced (its own log density + what absorbs
        at its border), the new value by sampleLogCategorical (utils.py:170-180; draw 2), always accepted (draw 1 is still
        drawn).  Returns (changed counts, mean node counts)."""
        step0, self.step = self.step, self.step + n_steps
        changed, nodes = np.zeros(self.P, dtype=np.int64), np.zeros(self.P)
        for p in range(self.P):
            for t in range(n_steps):
                u = oracle.uniforms(self.seed, self.p_base + p, step0 + t, 2)
                k = self.pick(u[0], t, block, scope)
                old = float(self.sites[k, p])
                xi = []
                cand = self.candidates[k]
                if not cand:
                    raise ValueError("gibbs: site %d cannot be enumerated" % k)
                for v in cand:
                    self.sites[k, p] = v
                    R = [0.0] * self.n_regs
                    xi.append(self.children(k, R, p) + self.run(self.site[k][2], R, p))
                mx = max(xi)
                es = [math.exp(x - mx) for x in xi] if mx > -math.inf else [0.0] * len(xi)
                den = sum(es)
                ps = [e / den for e in es] if den > 0 else [1.0 / len(xi)] * len(xi)
                if map:                                         # EnumerativeMAPOperator, egibbs.py:117-121: no draw
                    idx = [i for i, x in enumerate(xi) if x == mx][0]
                else:
                    idx = int(self._tape(p, step0 + t, 2)(3, ps))
                new = cand[idx]
                if not math.log(u[1]) < 0.0:
                    new = old
                self.sites[k, p] = new
                changed[p] += new != old
                nodes[p] += self.affected[k]
        return changed, nodes / max(n_steps, 1)

    def slice(self, n_steps, w, m, block=-1, scope=0, doubling=False):
        """StepOutSliceOperator (infer/slice_sample.py:39-139) or, doubling=True, DoublingSliceOperator (:148-196, m = the
        number of doublings p) under mixMH: draw 0 block choice, then the uniforms in call order.  Returns per-particle
        mean node counts."""
        step0, self.step = self.step, self.step + n_steps
        nodes = np.zeros(self.P)
        for p in range(self.P):
            for t in range(n_steps):
                tape = self._tape(p, step0 + t, 0)
                u_sel = tape(0, [0.0])
                k = self.pick(u_sel, t, block, scope)
                R = [0.0] * self.n_regs

                def f(x):
                    self.sites[k, p] = x
                    return self.children(k, R, p) + self.run(self.site[k][2], R, p)
                x0 = float(self.sites[k, p])
                logy = f(x0) + math.log(tape(0, [0.0]))
                L = x0 - w * tape(0, [0.0])
                Rr = L + w
                if doubling:                                     # :155-176
                    K = m
                    fl, fr = f(L), f(Rr)
                    while K > 0:
                        if logy >= fl and logy >= fr:
                            break
                        if math.isnan(fl) or math.isnan(fr):
                            break
                        dist = Rr - L
                        if tape(0, [0.0]) < 0.5:
                            L = L - dist
                            fl = f(L)
                        else:
                            Rr = Rr + dist
                            fr = f(Rr)
                        K -= 1
                else:
                    J = math.floor(m * tape(0, [0.0]))
                    K = (m - 1) - J
                    while J > 0:
                        fl = f(L)
                        if logy >= fl or math.isnan(fl):
                            break
                        L, J = L - w, J - 1
                    while K > 0:
                        fr = f(Rr)
                        if logy >= fr or math.isnan(fr):
                            break
                        Rr, K = Rr + w, K - 1

                def legal(x1, L, Rr):                            # :178-196 (stepping out: always legal, :146)
                    if not doubling:
                        return True
                    D = False
                    fr, fl = f(Rr), f(L)
                    while (Rr - L) > 1.1 * w:
                        M = (L + Rr) / 2.0
                        if (x0 < M and x1 >= M) or (x0 >= M and x1 < M):
                            D = True
                        if x1 < M:
                            Rr = M
                            fr = f(Rr)
                        else:
                            L = M
                            fl = f(L)
                        if D and logy >= fl and logy >= fr:
                            return False
                    return True
                x1 = x0
                for _ in range(10000):
                    xc = L + tape(0, [0.0]) * (Rr - L)
                    if logy <= f(xc) and legal(xc, L, Rr):
                        x1 = xc
                        break
                    if xc < x0:
                        L = xc
                    else:
                        Rr = xc
                if not math.log(tape(0, [0.0])) < 0.0:
                    x1 = x0
                self.sites[k, p] = x1
                nodes[p] += self.affected[k]
        return nodes / max(n_steps, 1)

    def subsampled_mh(self, n_steps, Nbatch, k0, eps, block=-1, scope=0):
        """subsampledMixMH + sequentialTest (infer/subsampled_mh.py:76-171) for a compiled program: draw 0 block choice,
        draw 1 the accept uniform, then the proposal; the minibatch order is the Feistel permutation of (seed, step)
        (the stand-in behind np_rng.permutation in tests/golden/make_program_chains.py); q = scipy.stats.t.cdf as in
        the reference.  Returns (accepted, mean sections visited, last t statistic, mean node counts) per particle."""
        import ctypes
        step0, self.step = self.step, self.step + n_steps
        acc, visited, tlast, nodes = np.zeros(self.P), np.zeros(self.P), np.zeros(self.P), np.zeros(self.P)
        for t in range(n_steps):
            perms = {}
            for p in range(self.P):
                u = oracle.uniforms(self.seed, self.p_base + p, step0 + t, 2)
                k = self.pick(u[0], t, block, scope)
                kind, N, off = self.ss[k]
                if kind == 2:
                    raise ValueError("subsampled_mh: site %d has no valid global border" % k)
                logu = math.log(u[1])
                nodes[p] += self.affected[self.S + k]
                old = float(self.sites[k, p])
                if kind == 0:                                   # no local sections: regular MH (:91-95)
                    R = [0.0] * self.n_regs
                    rho = self.children(k, R, p)
                    self.run(self.site[k][0], R, p, tape=self._tape(p, step0 + t, 2))
                    xi = self.children(k, R, p)
                    with np.errstate(invalid="ignore"):
                        alpha = np.float64(xi) - np.float64(rho)
                    if alpha > logu:
                        acc[p] += 1
                    else:
                        self.sites[k, p] = old
                    continue
                Rold, Rnew = [0.0] * self.n_regs, [0.0] * self.n_regs
                for g in self.site_groups[k]:
                    self.run(self.group[g][0], Rold, p)
                self.run(self.site[k][0], Rnew, p, tape=self._tape(p, step0 + t, 2))
                new = float(self.sites[k, p])
                for g in self.site_groups[k]:
                    self.run(self.group[g][0], Rnew, p)
                if N not in perms:
                    perm = np.empty(N, dtype=np.uint64)
                    oracle.lib().orc_feistel_perm(ctypes.c_uint64(self.seed), ctypes.c_uint32(step0 + t), ctypes.c_uint64(N),
                                                  perm.ctypes.data_as(ctypes.c_void_p))
                    perms[N] = perm
                perm = perms[N]

                def dllh(q):
                    g, r = self.roots[off + int(perm[q])]
                    if g >= 0:
                        _, row_sec, _, n_cols, doff = self.group[g]
                        row = self.data[doff + r * n_cols: doff + (r + 1) * n_cols]
                        return self.run(row_sec, Rnew, p, row=row) - self.run(row_sec, Rold, p, row=row)
                    self.sites[k, p] = old
                    lo = self.run(self.site[r][2], [0.0] * self.n_regs, p)
                    self.sites[k, p] = new
                    return self.run(self.site[r][2], [0.0] * self.n_regs, p) - lo
                accept, n, ts = sequential_test((logu - 0.0) / float(N), k0, Nbatch, N, eps, dllh)
                visited[p] += n
                if ts is not None:
                    tlast[p] = ts
                if accept:
                    acc[p] += 1
                else:
                    self.sites[k, p] = old
        return acc, visited / max(n_steps, 1), tlast, nodes / max(n_steps, 1)


def sequential_test(mu_0, k0, Nbatch, N, epsilon, fun_dllh):
    """infer/subsampled_mh.py:118-171 (returns the raw t statistic, None when none was computed)"""
    Nbatch, N = float(Nbatch), float(N)
    k = n = 0.0
    cum = cum2 = 0.0
    tstat = accept = None
    while n < N:
        n_end = min(n + Nbatch, N)
        for i in range(int(n), int(n_end)):
            d = fun_dllh(i)
            cum += d
            cum2 += d * d
        k += 1
        n = n_end
        mx, mx2 = cum / n_end, cum2 / n_end
        if k < k0 and n < N:
            continue
        if n == N:
            accept = mx >= mu_0
            break
        with np.errstate(invalid="ignore", divide="ignore"):
            sx = np.sqrt((1 - (n - 1) / (N - 1)) * (mx2 - mx * mx) / (n - 1))
        if sx == 0:
            q = 1 if mx >= mu_0 else 0
        else:
            tstat = (mx - mu_0) / sx
            q = scipy.stats.t.cdf(tstat, n - 1)
        if q < epsilon:
            accept = False
            break
        elif q > 1 - epsilon:
            accept = True
            break
    return bool(accept), n, tstat
