"""Model compiler: a fixed-structure program of `assume` / `observe` directives -> site programs for the
generic particle kernels (csrc/k_program.cu), SURVEY 8(f) N2.

What the reference does per transition -- `constructScaffold` (backend/lite/scaffold.py:149-173) from the
principal node, `extendCandidateScaffold` (:201-218: lookups and deterministic applications downstream are
resampled, random applications absorb: `canAbsorb`, psp.py:136-148, 459-461), `findBorder` (:260-264),
then `detachAndExtract` / `regenAndAttach` over that scaffold (detach.py:31-54, regen.py:49-72) -- is done
ONCE here, per random choice, for programs whose structure cannot change:

  expressions   numbers, booleans, symbols bound by earlier assumes, applications of
  deterministic SPs   add sub mul div pow sqrt exp log log1p expm1 logistic log_logistic logit negate abs min max
                      (venmath.py:101-332; add / mul are variadic left folds); `(array e ...)` and
                      `(vector_dot a b)` of two such arrays (vectors.py), lowered to scalar multiply-adds
  random SPs          normal gamma beta uniform_continuous expon flip bernoulli log_flip log_bernoulli
                      log_odds_flip log_odds_bernoulli            (continuous.py, discrete.py)

(no `if` on random predicates, no `mem`, no procedures: nothing that creates brush.)  Every unobserved random
application is a SITE = one block of the default scope (lite/trace.py:116-119).  For site k the compiler
emits, as straight-line code for a register machine,

  SIM_k        the site's parameters from its parents, then `simulate`                (regen.py:255-285 applyPSP)
  CHILD_k      the log densities of the random applications that absorb at k's border -- other sites and
               single observations -- with the deterministic nodes between recomputed  (regen.py:120-127 absorb)
  groups       observations that share one expression up to numeric literals are one GROUP: a prologue (the
               row-independent nodes), a row program, and a table with one row of literals + observed value per
               observation; a site lists the groups below it

and the number of nodes Lite's scaffold would hold (`numAffectedNodes`, scaffold.py:67-68: the site, the
lookup nodes of symbols bound to resampled nodes, the deterministic applications downstream), which is what
`resimulation_mh` reports (dispatch.py:109-120).  Single-site MH then is: rho = CHILD_k + groups under the old
value, SIM_k, xi = the same under the new value, accept iff log u < xi - rho (mh.py:84-89, 187-208; the prior
proposal contributes 0 to both).

`CompiledModel` plugs into B200TraceSet like the hand-described models of models.py and carries the program
as its `program()` for interchange.py, so device particles restore into real Lite traces.
"""
import ctypes

import numpy as np

from . import _lib
from . import interchange
from .exception import VentureException

# ---- the register machine (mirrored in csrc/k_program.cu and oracle/oracle.c) ----------------------------
OP = dict(
    CONST=0, SITE=1, DATA=2, ADD=3, SUB=4, MUL=5, DIV=6, NEG=7, EXP=8, LOG=9, SQRT=10, POW=11, LOGISTIC=12,
    ABS=13, MIN=14, MAX=15, LOG1P=16, EXPM1=17, LOG_LOGISTIC=18, LOGIT=19, STORE=20,
    COND=21, SELECT=22, MOV=23, ARGN=24, LOOKUP=25,
    # (dense numbering: the interpreter's switch is one jump table)
    SIM_NORMAL=26, SIM_GAMMA=27, SIM_BETA=28, SIM_UNIFORM=29, SIM_EXPON=30, SIM_FLIP=31, SIM_LOG_FLIP=32,
    SIM_LOG_ODDS_FLIP=33, SIM_CATEGORICAL=34,
    LD_NORMAL=35, LD_GAMMA=36, LD_BETA=37, LD_UNIFORM=38, LD_EXPON=39, LD_FLIP=40, LD_LOG_FLIP=41,
    LD_LOG_ODDS_FLIP=42, LD_CATEGORICAL=43)
INSTR_WORDS = 4          # op, dst, a, b
MAX_REGS = 128

DET = {"add": ("fold", "ADD"), "mul": ("fold", "MUL"), "sub": (2, "SUB"), "div": (2, "DIV"), "pow": (2, "POW"),
       "min": (2, "MIN"), "max": (2, "MAX"), "sqrt": (1, "SQRT"), "exp": (1, "EXP"), "log": (1, "LOG"),
       "log1p": (1, "LOG1P"), "expm1": (1, "EXPM1"), "logistic": (1, "LOGISTIC"), "log_logistic": (1, "LOG_LOGISTIC"),
       "logit": (1, "LOGIT"), "negate": (1, "NEG"), "abs": (1, "ABS"),
       # (biplex pred a b), conditionals.py:27-37: both branches are evaluated, no brush; two instructions: COND, SELECT
       "biplex": (3, "BIPLEX")}
RANDOM = {"normal": (2, "NORMAL"), "gamma": (2, "GAMMA"), "beta": (2, "BETA"), "uniform_continuous": (2, "UNIFORM"),
          "expon": (1, "EXPON"), "flip": (1, "FLIP"), "bernoulli": (1, "FLIP"), "log_flip": (1, "LOG_FLIP"),
          "log_bernoulli": (1, "LOG_FLIP"), "log_odds_flip": (1, "LOG_ODDS_FLIP"), "log_odds_bernoulli": (1, "LOG_ODDS_FLIP"),
          # (categorical <simplex or array> [<array of literal objects>]), discrete.py:208-246; values default to 0..n-1
          "categorical": ("cat", "CATEGORICAL")}
BOOLEAN_VALUED = ("FLIP", "LOG_FLIP", "LOG_ODDS_FLIP")


def _tag_key(e):
    """scope / block operand of `tag` or of an inference command -> a hashable python value: symbols (quoted or not) as
    str, numbers as float"""
    e = _plain(e) if not isinstance(e, (str, float, int, bool, list)) else e
    if isinstance(e, list):
        if len(e) == 2 and e[0] == "quote":
            return _tag_key(e[1])
        raise VentureException("invalid_argument", "scope / block operands of tag must be literals, not %r" % (e,))
    if isinstance(e, bool) or isinstance(e, str):
        return e
    return float(e)


def _plain(e):
    """stack-dict expression / VentureValue -> nested python: str symbols, float numbers, bools, lists"""
    if hasattr(e, "asStackDict"):
        e = e.asStackDict()
    if isinstance(e, dict):
        t, v = e.get("type"), e.get("value")
        if t == "symbol":
            return str(v)
        if t == "boolean":
            return bool(v)
        if isinstance(v, (int, float, np.integer, np.floating)) and not isinstance(v, bool):
            return float(v)
        raise VentureException("invalid_argument", "the model compiler has no value of type %r" % (t,))
    if isinstance(e, (list, tuple)):
        return [_plain(x) for x in e]
    if isinstance(e, bool):
        return e
    if isinstance(e, (int, float, np.integer, np.floating)):
        return float(e)
    if isinstance(e, str):
        return e
    raise VentureException("invalid_argument", "cannot compile %r" % (e,))


class Node(object):
    __slots__ = ("kind", "op", "args", "value", "site", "hole", "deps", "rowvar", "lookups", "counted")

    def __init__(self, kind, op=None, args=(), value=None):
        self.kind, self.op, self.args, self.value = kind, op, tuple(args), value   # kind: const | hole | det | random
        self.site = None       # site index for an unobserved random application
        self.hole = None       # column of the group's table for a literal of an observation template
        self.deps = frozenset()   # sites this node depends on (through deterministic paths AND as a random node's parameters)
        self.rowvar = False
        self.lookups = 0       # lookup nodes (symbol occurrences) whose source is this node
        self.counted = True    # False for the inner steps of a variadic add / mul: ONE application node in Lite


class Group(object):
    """observations sharing one expression up to numeric literals"""

    def __init__(self, template, root, n_holes):
        self.template, self.root, self.n_holes = template, root, n_holes
        self.rows = []         # per observation: literals..., observed value


class Compiler(object):
    def __init__(self):
        self.nodes = []
        self.env = {}          # symbol -> node id
        self.sites = []        # node ids of the sites, in evaluation order
        self.site_names = []
        self.single_obs = []   # (node id, value) of observations compiled one by one
        self.groups = {}       # template key -> Group
        self.group_list = []
        self.directives = []   # the program, for interchange
        self.directive_sites = []   # per directive: the sites its evaluation creates (none for observes)
        # Lite's child lists, for findGlobalBorder (infer/subsampled_mh.py:326-368): node -> its consumers in
        # creation order, [consumer node, through a lookup node?, group or None, row]
        self.uses = {}
        # scopes made by `(tag scope block expr)`: scope -> block -> sites created inside expr, both in order of first
        # appearance (trace.registerRandomChoiceInScope, lite/trace.py:121-127; regen.py:343-348)
        self.scopes = {}
        self.aliased = set()   # nodes some `(assume y x)` renames: their lookups hang off the alias, not off the node
        self._template_uses = None

    # -- building the node graph -------------------------------------------------------------------------
    def _new(self, node):
        self.nodes.append(node)
        return len(self.nodes) - 1

    def _expr(self, e, holes=None, observed_root=False):
        """-> node id.  holes: list collecting the numeric literals of an observation template, or None"""
        if isinstance(e, bool):
            return self._new(Node("const", value=1.0 if e else 0.0))
        if isinstance(e, float):
            if holes is not None:
                n = Node("hole")
                n.hole, n.rowvar = len(holes), True
                holes.append(e)
                return self._new(n)
            return self._new(Node("const", value=e))
        if isinstance(e, str):
            if e not in self.env:
                raise VentureException("evaluation", "Cannot find symbol '%s'" % (e,))
            src = self.env[e]
            self.nodes[src].lookups += 1
            return src
        if not e:
            raise VentureException("invalid_argument", "empty application")
        head = e[0]
        if not isinstance(head, str):
            raise VentureException("invalid_argument", "the model compiler needs a builtin SP in operator position, not %r" % (head,))
        if head in self.env:
            raise VentureException("invalid_argument", "procedure application (%s ...) is outside the fixed-structure subset" % (head,))
        if head == "tag":
            # scope.py:25-52: returns its third operand unchanged (ONE deterministic node in the scaffolds) and registers the
            # random choices made inside it in scope / block
            if len(e) != 4 or holes is not None or not isinstance(e[3], list):
                raise VentureException("invalid_argument", "(tag scope block (application ...)) outside an observation is the compiled form")
            scope, block = _tag_key(e[1]), _tag_key(e[2])
            if scope == "default":
                raise VentureException("invalid_argument", "the default scope cannot be tagged into")
            first = len(self.sites)
            inner = self._expr(e[3])
            blocks = self.scopes.setdefault(scope, {})
            blocks.setdefault(block, []).extend(range(first, len(self.sites)))
            return self._record_uses(self._finish(Node("det", "TAG", [inner])), [inner], [False])
        args = [self._expr(a, holes) for a in e[1:]]
        via = [isinstance(a, str) for a in e[1:]]
        if head == "lookup":
            # (lookup <array> <index>), one deterministic application; the index is a value computed on the device
            if len(args) != 2 or self.nodes[args[0]].op != "ARRAY":
                raise VentureException("invalid_argument", "lookup takes an array built by (array ...) / (simplex ...) and an index in the compiled subset")
            return self._record_uses(self._finish(Node("det", "LOOKUP", args)), args, via)
        if head in ("array", "simplex"):
            # `(array e_1 ... e_n)` (vectors.py): ONE deterministic application whose value is the tuple of its operands;
            # it owns no register -- vector_dot reads the operands' registers
            return self._record_uses(self._finish(Node("det", "ARRAY", args)), args, via)
        if head == "vector_dot":
            if len(args) != 2 or any(self.nodes[a].op != "ARRAY" for a in args):
                raise VentureException("invalid_argument", "vector_dot takes two arrays built by (array ...) in the compiled subset")
            if len(self.nodes[args[0]].args) != len(self.nodes[args[1]].args):
                raise VentureException("invalid_argument", "vector_dot: the arrays differ in length")
            return self._record_uses(self._finish(Node("det", "VDOT", args)), args, via)
        if head in DET:
            arity, op = DET[head]
            if arity == "fold":
                if not args:
                    return self._new(Node("const", value=0.0 if op == "ADD" else 1.0))
                acc = args[0]
                for i, a in enumerate(args[1:]):
                    acc = self._finish(Node("det", op, (acc, a)))
                    self.nodes[acc].counted = i == len(args) - 2
                if len(args) > 1:
                    self._record_uses(acc, args, via)        # ONE application node in Lite: every operand's consumer
                return acc
            if len(args) != arity:
                raise VentureException("invalid_argument", "%s takes %d argument(s)" % (head, arity))
            return self._record_uses(self._finish(Node("det", op, args)), args, via)
        if head in RANDOM:
            arity, op = RANDOM[head]
            if arity == "cat":
                if len(args) not in (1, 2) or any(self.nodes[a].op != "ARRAY" for a in args):
                    raise VentureException("invalid_argument", "categorical takes (simplex ...) or (array ...) weights and optionally an array of objects")
                if len(args) == 2:
                    objs = [self.nodes[o] for o in self.nodes[args[1]].args]
                    if len(objs) != len(self.nodes[args[0]].args) or any(o.kind != "const" for o in objs):
                        raise VentureException("invalid_argument", "categorical: the objects must be as many literal numbers as there are weights")
                if not 1 <= len(self.nodes[args[0]].args) <= 32:
                    raise VentureException("invalid_argument", "categorical: 1 to 32 outcomes are compiled")
                arity = len(args)
            if op == "FLIP" and not args:
                args = [self._new(Node("const", value=0.5))]           # (flip) = (flip 0.5), discrete.py:98-102
            if len(args) != arity:
                raise VentureException("invalid_argument", "%s takes %d argument(s)" % (head, arity))
            nid = self._record_uses(self._finish(Node("random", op, args)), args, via)
            if not observed_root:
                if holes is not None:
                    raise VentureException("invalid_argument", "a latent random choice inside an observed expression is outside the fixed-structure subset")
                self.nodes[nid].site = len(self.sites)
                self.sites.append(nid)
                self.site_names.append(None)
            return nid
        raise VentureException("invalid_argument", "SP %r is outside the model compiler's subset" % (head,))

    def _record_uses(self, consumer, args, via):
        for a, lookup in zip(args, via):
            use = [consumer, lookup, None, 0]
            self.uses.setdefault(a, []).append(use)
            if self._template_uses is not None:
                self._template_uses.append((a, use))
        return consumer

    def _finish(self, node):
        deps, rowvar = set(), False
        for a in node.args:
            an = self.nodes[a]
            rowvar = rowvar or an.rowvar
            if an.kind == "random":
                if an.site is None:
                    raise VentureException("invalid_argument", "an observed value cannot be a parent in the fixed-structure subset")
                deps.add(an.site)
            else:
                deps |= an.deps
        node.deps, node.rowvar = frozenset(deps), rowvar
        return self._new(node)

    # -- directives -----------------------------------------------------------------------------------------
    def assume(self, name, exp):
        exp = _plain(exp)
        first = len(self.sites)
        nid = self._expr(exp)
        n = self.nodes[nid]
        while n.op == "TAG":
            n = self.nodes[n.args[0]]
        if not isinstance(exp, str) and n.kind == "random":
            self.site_names[n.site] = name          # (an alias's root is the lookup counted in _expr)
        if isinstance(exp, str):
            self.aliased.add(nid)
        self.env[name] = nid
        self.directives.append(interchange.define(name, exp))
        self.directive_sites.append(list(range(first, len(self.sites))))
        return nid

    def observe(self, exp, value):
        exp = _plain(exp)
        v = _plain(value)
        v = (1.0 if v else 0.0) if isinstance(v, bool) else float(v)
        if not (isinstance(exp, list) and exp and exp[0] in RANDOM):
            raise VentureException("invalid_argument", "only applications of random SPs can be observed (psp.py:459-461), not %r" % (exp,))
        key = self._template_key(exp)
        g = self.groups.get(key)
        if g is None:
            holes = []
            start = len(self.nodes)
            self._template_uses = []
            root = self._expr_observed(exp, holes)
            g = Group(key, root, len(holes))
            g.own = range(start, len(self.nodes))      # the template's own nodes: one copy per observation in Lite
            g.symbol_uses = [(a, use[0]) for a, use in self._template_uses if use[1]]
            for _, use in self._template_uses:
                use[2] = g
            self._template_uses = None
            self.groups[key] = g
            self.group_list.append(g)
            g.rows.append(holes + [v])
        else:
            lits = []
            self._literals(exp, lits)
            g.rows.append(lits + [v])
            self._count_template_lookups(exp)
            for a, consumer in g.symbol_uses:          # this observation's own lookup nodes, in evaluation order
                self.uses[a].append([consumer, True, g, len(g.rows) - 1])
        boolean = RANDOM[exp[0]][1] in BOOLEAN_VALUED
        self.directives.append(interchange.observe(exp, interchange.boolean(v != 0.0) if boolean else interchange.number(v)))
        self.directive_sites.append([])

    def _expr_observed(self, exp, holes):
        arity, op = RANDOM[exp[0]]
        args = [self._expr(a, holes) for a in exp[1:]]
        via = [isinstance(a, str) for a in exp[1:]]
        if op == "FLIP" and not args:
            args, via = [self._new(Node("const", value=0.5))], [False]
        if len(args) != arity:
            raise VentureException("invalid_argument", "%s takes %d argument(s)" % (exp[0], arity))
        return self._record_uses(self._finish(Node("random", op, args)), args, via)

    def _template_key(self, e):
        if isinstance(e, bool):
            return ("b", e)
        if isinstance(e, float):
            return "#"
        if isinstance(e, str):
            return ("s", e)
        return tuple(self._template_key(x) for x in e)

    def _literals(self, e, out):
        if isinstance(e, float) and not isinstance(e, bool):
            out.append(e)
        elif isinstance(e, list):
            for x in e[1:]:
                self._literals(x, out)

    def _count_template_lookups(self, e):
        if isinstance(e, str):
            self.nodes[self.env[e]].lookups += 1
        elif isinstance(e, list):
            for x in e[1:]:
                self._count_template_lookups(x)

    # -- code generation --------------------------------------------------------------------------------------
    def compile(self):
        return Program(self)


class _Code(object):
    """one straight-line section; registers are allocated per section from a shared pool of node -> register"""

    def __init__(self, prog):
        self.prog, self.words, self.done = prog, [], set()

    def emit(self, op, dst=0, a=0, b=0):
        self.words += [OP[op], dst, a, b]

    def reg(self, nid):
        return self.prog.reg_of(nid)

    def need(self, nid, fresh_sites=()):
        """make node nid's value available in its register (recursively); sites are loaded from memory"""
        if nid in self.done:
            return self.reg(nid)
        n = self.prog.c.nodes[nid]
        if n.op == "TAG":                            # the operand's register under a second name
            r = self.need(n.args[0])
            self.prog._regs[nid] = r
            self.done.add(nid)
            return r
        r = self.reg(nid)
        if n.kind == "const":
            self.emit("CONST", r, self.prog.const(n.value))
        elif n.kind == "hole":
            self.emit("DATA", r, n.hole)
        elif n.kind == "random":
            self.emit("SITE", r, n.site)
        elif n.op == "BIPLEX":
            regs = [self.need(a) for a in n.args]
            self.emit("COND", 0, regs[0])
            self.emit("SELECT", r, regs[1], regs[2])
        elif n.op == "LOOKUP":
            arr = self.prog.c.nodes[n.args[0]]
            self.need(n.args[0])
            idx = self.need(n.args[1])
            base = self.block([self.reg(a) for a in arr.args], ("lookup", nid))
            self.emit("ARGN", 0, len(arr.args))
            self.emit("LOOKUP", r, base, idx)
        elif n.op == "ARRAY":
            for a in n.args:
                self.need(a)
        elif n.op == "VDOT":                         # left fold of the products (vectors.py vector_dot = numpy.dot)
            xs, ys = self.prog.c.nodes[n.args[0]], self.prog.c.nodes[n.args[1]]
            self.need(n.args[0]), self.need(n.args[1])
            pairs = [(self.reg(a), self.reg(b)) for a, b in zip(xs.args, ys.args)]
            if not pairs:
                self.emit("CONST", r, self.prog.const(0.0))
            else:
                self.emit("MUL", r, pairs[0][0], pairs[0][1])
                tmp = self.prog.reg_of(("vdot-tmp", nid))
                for ra, rb in pairs[1:]:
                    self.emit("MUL", tmp, ra, rb)
                    self.emit("ADD", r, r, tmp)
        else:
            regs = [self.need(a) for a in n.args]
            self.emit(n.op, r, regs[0], regs[1] if len(regs) > 1 else 0)
        self.done.add(nid)
        return r

    def block(self, regs, key):
        """copies of `regs` in CONSECUTIVE registers (the variadic instructions address a base register); -> base"""
        dst = [self.prog.reg_of((key, j)) for j in range(len(regs))]
        assert dst == list(range(dst[0], dst[0] + len(dst))) if dst else True
        for d, r in zip(dst, regs):
            self.emit("MOV", d, r)
        return dst[0] if dst else 0

    def categorical_block(self, nid):
        """weights then objects of a categorical in consecutive registers; -> (base, n)"""
        n = self.prog.c.nodes[nid]
        nodes = self.prog.c.nodes
        ws = nodes[n.args[0]].args
        self.need(n.args[0])
        regs = [self.reg(a) for a in ws]
        if len(n.args) == 2:
            self.need(n.args[1])
            regs += [self.reg(a) for a in nodes[n.args[1]].args]
        else:
            for j in range(len(ws)):
                c = self.prog.reg_of(("cat-obj", nid, j))
                self.emit("CONST", c, self.prog.const(float(j)))
                regs.append(c)
        return self.block(regs, ("cat", nid)), len(ws)

    def log_density(self, nid, value_reg=None):
        n = self.prog.c.nodes[nid]
        if n.op == "CATEGORICAL":
            base, k = self.categorical_block(nid)
            x = value_reg if value_reg is not None else self.need(nid)
            self.emit("ARGN", 0, k)
            self.emit("LD_CATEGORICAL", x, base)
            return
        regs = [self.need(a) for a in n.args]
        x = value_reg if value_reg is not None else self.need(nid)
        self.emit("LD_" + n.op, x, regs[0], regs[1] if len(regs) > 1 else 0)

    def simulate(self, nid):
        n = self.prog.c.nodes[nid]
        if n.op == "CATEGORICAL":
            base, k = self.categorical_block(nid)
            r = self.reg(nid)
            self.emit("ARGN", 0, k)
            self.emit("SIM_CATEGORICAL", r, base)
            self.emit("STORE", r, n.site)
            self.done.add(nid)
            return
        regs = [self.need(a) for a in n.args]
        r = self.reg(nid)
        self.emit("SIM_" + n.op, r, regs[0], regs[1] if len(regs) > 1 else 0)
        self.emit("STORE", r, n.site)
        self.done.add(nid)


class Program(object):
    """The compiled blob: int32 code words, double constants, per-group tables, section directory."""

    def __init__(self, c):
        self.c = c
        self._regs, self._consts, self._const_ix = {}, [], {}
        nodes = c.nodes
        S = len(c.sites)
        if S == 0:
            raise VentureException("invalid_argument", "the program has no unobserved random choice: nothing to infer")
        # children of every site: random nodes (sites / group roots) whose parameters depend on it
        site_children = [[] for _ in range(S)]
        for nid in c.sites:
            for s in nodes[nid].deps:
                site_children[s].append(nid)
        site_groups = [[] for _ in range(S)]
        for gi, g in enumerate(c.group_list):
            for s in nodes[g.root].deps:
                site_groups[s].append(gi)
        sections = []

        def add(code):
            sections.append(code.words)
            return len(sections) - 1
        # 0: PRIOR -- every site in evaluation order
        code = _Code(self)
        for nid in c.sites:
            code.simulate(nid)
        self.sec_prior = add(code)
        # 1: JOINT -- log density of every site given its parents
        code = _Code(self)
        for nid in c.sites:
            code.log_density(nid)
        self.sec_joint = add(code)
        self.sec_sim, self.sec_child, self.sec_own = [], [], []
        for k, nid in enumerate(c.sites):
            code = _Code(self)
            code.simulate(nid)
            self.sec_sim.append(add(code))
            code = _Code(self)
            for ch in site_children[k]:
                code.log_density(ch)
            self.sec_child.append(add(code))
            code = _Code(self)
            code.log_density(nid)
            self.sec_own.append(add(code))
        self.sec_group_pro, self.sec_group_row = [], []
        for g in c.group_list:
            pro = _Code(self)
            root = nodes[g.root]
            for nid in self._invariant_frontier(g.root):
                pro.need(nid)
            self.sec_group_pro.append(add(pro))
            row = _Code(self)
            row.done = set(pro.done)
            vreg = self.reg_of(("value", id(g)))
            row.emit("DATA", vreg, g.n_holes)
            row.log_density(g.root, value_reg=vreg)
            self.sec_group_row.append(add(row))
            assert root.kind == "random"
        self.site_groups = site_groups
        self.n_sites, self.n_groups = S, len(c.group_list)
        self.n_regs = len(self._regs)
        if self.n_regs > MAX_REGS:
            raise VentureException("invalid_argument", "the program needs %d registers; the generic kernel has %d" % (self.n_regs, MAX_REGS))
        self.sections = sections
        # numAffectedNodes per site (scaffold.py:67-68): the site + lookups of resampled nodes + deterministic applications downstream
        self.affected = []
        own = {}
        for g in c.group_list:
            for nid in g.own:
                own[nid] = len(g.rows)
        for k in range(S):
            cnt = 1 + nodes[c.sites[k]].lookups
            for nid, n in enumerate(nodes):
                if n.kind == "det" and k in n.deps:          # deps stop at random nodes: these are k's deterministic descendants
                    cnt += (own.get(nid, 1) if n.counted else 0) + n.lookups
            self.affected.append(float(cnt))
        self.subsampled = [self._subsample_site(k) for k in range(S)]
        # what enumerative gibbs / emap try at a site (PSP.enumerateValues): [true, false] for the flip family
        # (discrete.py:78-86, 121-125, 156-164), the objects of a categorical (:229-232); nothing for continuous SPs
        self.candidates = []
        for nid in c.sites:
            n = nodes[nid]
            if n.op in BOOLEAN_VALUED:
                self.candidates.append([1.0, 0.0])
            elif n.op == "CATEGORICAL":
                k = len(nodes[n.args[0]].args)
                self.candidates.append([float(nodes[o].value) for o in nodes[n.args[1]].args] if len(n.args) == 2 else [float(j) for j in range(k)])
            else:
                self.candidates.append([])
        # scope 0 = default (one block per site, lite/trace.py:116-119), then the tagged scopes; a block is one site here
        self.scope_names, self.scope_sites, self.scope_blocks = ["default"], [list(range(S))], [dict((k, k) for k in range(S))]
        for name, blocks in c.scopes.items():
            sites, index = [], {}
            for key, members in blocks.items():
                if len(members) != 1:
                    raise VentureException("invalid_argument", "scope %r block %r holds %d random choices: blocks of one random "
                                           "choice are the compiled subset (single-site proposals)" % (name, key, len(members)))
                index[key] = len(sites)
                sites.append(members[0])
            self.scope_names.append(name), self.scope_sites.append(sites), self.scope_blocks.append(index)

    # -- subsampled MH: the global border and the local sections of a site (infer/subsampled_mh.py:326-368, 186-211) --
    def _subsample_site(self, k):
        """-> dict(kind, reason, roots, nodes).  kind 0: no global border -- subsampledMixMH is regular MH (:91-95);
        1: border found, roots = Lite's `trace.childrenAt(globalBorder[0])` in creation order, each one local
        section given as (group, row) or (-1, child site); 2: a structure Lite rejects or mis-handles
        (SubsampledScaffoldError / checkApplicability's NotApplicable warning) -- refused at the call.
        nodes: numAffectedNodes of the scaffold cut at the border (hardBorder, scaffold.py:149-173)."""
        c = self.c
        nodes = c.nodes
        plain = dict(kind=0, reason="", roots=[], nodes=self.affected[k])

        def refuse(why):
            return dict(kind=2, reason=why, roots=[], nodes=0.0)
        n, count, random_on_path = c.sites[k], 1, False
        while True:                                   # P -> single path -> the first node with several children
            if n in c.aliased:
                return refuse("an alias of a node between the principal node and its border is outside the compiled subset")
            uses = c.uses.get(n, [])
            if not uses:
                return plain
            if len(uses) > 1:
                break
            consumer, lookup = uses[0][0], uses[0][1]
            count += 1 if lookup else 0
            cn = nodes[consumer]
            if cn.kind == "random":
                if cn.site is None:
                    return plain                      # an observed application: it absorbs and has no children
                random_on_path = True
            else:
                count += 1
            n = consumer
        if random_on_path:                            # sampleGlobalIndex :177-180 raises: the border is not resampling
            return refuse("Invalid global border (a random choice between the principal node and the border absorbs)")
        border, roots, seen = n, [], set()
        for consumer, lookup, g, row in c.uses[border]:
            if g is not None:
                if self._template_entries(g, k) != [border]:
                    return refuse("local sections overlap: an observation reads the principal node along two paths")
                roots.append((c.group_list.index(g), row))
                continue
            m = consumer
            while nodes[m].kind == "det":             # the local section: deterministic applications down to one absorber
                u = c.uses.get(m, [])
                if len(u) != 1 or m in c.aliased or u[0][2] is not None:
                    return refuse("a local section that fans out is outside the compiled subset")
                m = u[0][0]
            if nodes[m].site is None or m in seen:
                return refuse("local sections overlap")
            if k in nodes[m].deps and sum(1 for a in self._entries_of(m, k)) != 1:
                return refuse("local sections overlap: a child reads the principal node along two paths")
            seen.add(m)
            roots.append((-1, nodes[m].site))
        return dict(kind=1, reason="", roots=roots, nodes=float(count))

    def _depends(self, a, k):
        n = self.c.nodes[a]
        return n.site == k if n.kind == "random" else k in n.deps

    def _entries_of(self, root, k):
        """the occurrences through which site k reaches random node `root`'s parameters (anonymous nodes expanded)"""
        c = self.c
        named = set(c.env.values())
        out = []

        def walk(nid):
            for a in c.nodes[nid].args:
                if not self._depends(a, k):
                    continue
                if a in named or c.nodes[a].kind == "random":
                    out.append(a)
                else:
                    walk(a)
        walk(root)
        return out

    def _template_entries(self, g, k):
        return self._entries_of(g.root, k)

    def _invariant_frontier(self, root):
        """row-independent nodes a group's row program reads: computed once in the group prologue"""
        out = []

        def walk(nid):
            n = self.c.nodes[nid]
            if not n.rowvar and nid != root:
                out.append(nid)
                return
            for a in n.args:
                walk(a)
        walk(root)
        return out

    def reg_of(self, key):
        if key not in self._regs:
            self._regs[key] = len(self._regs)
        return self._regs[key]

    def const(self, v):
        key = np.float64(v).tobytes()
        if key not in self._const_ix:
            self._const_ix[key] = len(self._consts)
            self._consts.append(float(v))
        return self._const_ix[key]

    # -- the flat blob handed to vk_create_program (layout documented in include/venture_b200.h) ------------------
    def blob(self):
        code, directory = [], []
        for words in self.sections:
            directory += [len(code) // INSTR_WORDS, len(words) // INSTR_WORDS]
            code += words
        S, G = self.n_sites, self.n_groups
        meta = [S, G, self.n_regs, len(self.sections), self.sec_prior, self.sec_joint]
        for k in range(S):
            meta += [self.sec_sim[k], self.sec_child[k], self.sec_own[k], len(self.site_groups[k])]
        for k in range(S):
            meta += self.site_groups[k]
        data, off = [], 0
        for gi, g in enumerate(self.c.group_list):
            rows = np.asarray(g.rows, dtype=np.float64).reshape(len(g.rows), g.n_holes + 1)
            meta += [self.sec_group_pro[gi], self.sec_group_row[gi], rows.shape[0], rows.shape[1], off]
            data.append(rows.ravel())
            off += rows.size
        meta += directory
        # subsampled MH: per site (kind, number of local sections, offset into the root table), then the roots as pairs
        roots = []
        for k in range(S):
            ss = self.subsampled[k]
            meta += [ss["kind"], len(ss["roots"]), len(roots) // 2]
            for a, b in ss["roots"]:
                roots += [a, b]
        meta += [len(roots) // 2] + roots
        # scopes: count; per scope (blocks, offset into the site list); the blocks' sites
        meta += [len(self.scope_sites)]
        off = 0
        for sites in self.scope_sites:
            meta += [len(sites), off]
            off += len(sites)
        for sites in self.scope_sites:
            meta += sites
        # enumeration candidates: per site (count, offset), then indices into the constants
        off = 0
        for cand in self.candidates:
            meta += [len(cand), off]
            off += len(cand)
        for cand in self.candidates:
            meta += [self.const(v) for v in cand]
        return (np.asarray(code, dtype=np.int32), np.asarray(self._consts or [0.0], dtype=np.float64),
                np.asarray(meta, dtype=np.int32), np.concatenate(data) if data else np.zeros(1),
                np.asarray(self.affected + [ss["nodes"] for ss in self.subsampled], dtype=np.float64))


# ---- the model object B200TraceSet takes -----------------------------------------------------------------------
class CompiledModel(object):
    """A compiled program as a model of models.py: `B200TraceSet(CompiledModel(directives), P)`.
    directives: [("assume", name, exp) | ("observe", exp, value)] with expressions as nested python lists
    (`["normal", "mu", 1.0]`), stack dicts, or the host Venture's values."""
    kind = 5

    def __init__(self, directives):
        c = Compiler()
        for d in directives:
            if d[0] in ("assume", "define"):
                c.assume(d[1], d[2])
            elif d[0] == "observe":
                c.observe(d[1], d[2])
            else:
                raise VentureException("invalid_argument", "unknown directive %r" % (d[0],))
        self.compiler = c
        self.prog = c.compile()
        self.site_names = list(c.site_names)
        self.scopes = dict((name, i) for i, name in enumerate(self.prog.scope_names))       # B200TraceSet._scope_block
        self.scope_blocks = dict((i, b) for i, b in enumerate(self.prog.scope_blocks) if i > 0)
        self.n_observations = sum(len(g.rows) for g in c.group_list)
        self.source = list(directives)

    def extended(self, more):
        return CompiledModel(self.source + list(more))

    def _create(self, n_particles, seed, device):
        code, consts, meta, data, affected = self.prog.blob()
        h = ctypes.c_void_p()
        _lib.call("vk_create_program", n_particles, seed, device, _lib.ptr(code), code.size, _lib.ptr(consts), consts.size,
                  _lib.ptr(meta), meta.size, _lib.ptr(data), data.size, _lib.ptr(affected), ctypes.byref(h))
        return h

    def _observe(self, h, *args):
        raise VentureException("invalid_argument", "a compiled program holds its observations: extend the directive list "
                                                   "(CompiledModel.extended) instead of observing into the batch")

    # -- subsampled MH ------------------------------------------------------------------------------------------
    def subsampled_applicability(self, block):
        """SubsampledBlockScaffoldIndexer.checkApplicability (infer/subsampled_mh.py:222-262) from the compiler's
        analysis: block "one" walks every block and stops at the first that fails.  -> (ok, warning class name, message)"""
        sites = range(self.prog.n_sites) if block in ("one", "each", -1, -3) else [int(block)]
        for k in sites:
            ss = self.prog.subsampled[k]
            if ss["kind"] == 0:
                return False, "SubsampledScaffoldNotEffectiveWarning", "No global border found. subsampled_mh will behave the same as mh."
            if ss["kind"] == 2:
                return False, "SubsampledScaffoldNotApplicableWarning", ss["reason"]
        return True, None, ""

    def subsampled_global_nodes(self, block):
        """what SubsampledMHOperator.makeConsistent returns for a fresh operator (:411-441): the node counts of the global
        sections that have a border, summed over the blocks"""
        sites = range(self.prog.n_sites) if block in ("one", "each", -1, -3) else [int(block)]
        return float(sum(self.prog.subsampled[k]["nodes"] for k in sites if self.prog.subsampled[k]["kind"] == 1))

    # -- interchange.py ---------------------------------------------------------------------------------------
    def program(self, obs):
        return list(self.compiler.directives)

    def _random_order(self):
        """the random applications in evaluation order: ('site', k, boolean) | ('obs', value)"""
        out = []
        c = self.compiler
        for d, new_sites in zip(c.directives, c.directive_sites):
            for k in new_sites:
                out.append(("site", k, c.nodes[c.sites[k]].op in BOOLEAN_VALUED))
            if d[0] == "observe":
                out.append(("obs", d[2]))
        return out

    def random_values(self, obs, sites):
        out = []
        for item in self._random_order():
            if item[0] == "site":
                out.append(interchange.boolean(sites[item[1]] != 0.0) if item[2] else interchange.number(sites[item[1]]))
            else:
                out.append(item[1])
        return out

    def sites_from_random_values(self, obs, values):
        sites = np.zeros(self.prog.n_sites)
        for item, v in zip(self._random_order(), values):
            if item[0] == "site":
                sites[item[1]] = float(interchange.unbox(v))
        return sites
