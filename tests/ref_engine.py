"""TEST INFRASTRUCTURE: the reference's own control plane -- `venture.engine.engine.Engine`, its `Infer`
object, the inference SPs of backend/lite/inference_sps.py, the macro system (`do`, `let`, ...) and the whole
inference prelude -- imported from oracle/_ref (the py3 rendering of the reference, built by
oracle/lite_py3/build_lite_py3.py; present wherever /root/reference was present at build() time), so that
tests can run the reference's inference programs UNCHANGED against a backend of their choice: Lite itself, or
the B200 trace set (venturecxx_b200/backend.py).

The reference's parser needs a generated grammar module that is not built here; `read()` below is a small
s-expression reader that produces exactly what ChurchPrimeParser.parse_expression returns for the subset of
syntax the inference prelude and the test programs use (parse.py:171-224: symbols, the operator table,
true / false, numbers, strings, quote), and `desugar()` is the reference's own macro expansion followed by
core_sivm._modify_expression -- the pipeline of engine.py:156-165.
"""
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref")


def available():
    return os.path.isdir(os.path.join(REF, "venture", "engine")) and os.path.isdir(os.path.join(REF, "venture", "sivm"))


def _import():
    if REF not in sys.path:
        sys.path.insert(0, REF)
    import warnings
    warnings.simplefilter("ignore")


_OPERATORS = {"+": "add", "-": "sub", "*": "mul", "/": "div", "<": "lt", ">": "gt", "<=": "lte", ">=": "gte",
              "=": "eq", "!=": "neq", "<-": "<-"}
_TOKEN = re.compile(r"""\s*(?:;[^\n]*\n\s*)*([()']|"(?:[^"\\]|\\.)*"|[^\s()';]+)""")
_NUMBER = re.compile(r"^[+-]?(\d+\.?\d*([eE][+-]?\d+)?|\.\d+([eE][+-]?\d+)?)$")


def read(text):
    """church-prime expression text -> the stack-dict expression the reference's parser returns"""
    _import()
    import venture.value.dicts as val
    tokens = _TOKEN.findall(text)
    pos = [0]

    def atom(tok):
        if tok.startswith('"'):
            return val.string(bytes(tok[1:-1], "utf-8").decode("unicode_escape"))
        if tok == "true":
            return val.boolean(True)
        if tok == "false":
            return val.boolean(False)
        if _NUMBER.match(tok):
            return val.number(float(tok) if re.search(r"[.eE]", tok) else int(tok))
        return val.symbol(_OPERATORS.get(tok, tok))

    def parse():
        tok = tokens[pos[0]]
        pos[0] += 1
        if tok == "(":
            out = []
            while tokens[pos[0]] != ")":
                out.append(parse())
            pos[0] += 1
            return out
        if tok == "'":
            return val.quote(parse())
        if tok == ")":
            raise ValueError("unbalanced ) in %r" % (text,))
        return atom(tok)

    exp = parse()
    if pos[0] != len(tokens):
        raise ValueError("trailing tokens in %r" % (text,))
    return exp


def desugar(exp):
    """engine.py:163: _modify_expression(desugar_expression(parsed))"""
    _import()
    import venture.sivm.macro  # noqa: F401  (registers the macros)
    from venture.sivm.macro_system import desugar_expression
    from venture.sivm.core_sivm import _modify_expression
    return _listify(_modify_expression(desugar_expression(exp)))


def _listify(e):
    if isinstance(e, (list, tuple, map)):
        return [_listify(x) for x in e]
    return e


def program(text):
    """inference program text -> what the sivm hands to engine.infer"""
    return desugar(read(text))


def _install_prelude_reader():
    """engine.py:150-165 computes the prelude by parsing its forms; route that through read()/desugar()"""
    import venture.engine.engine as eng
    import venture.engine.inference_prelude as prelude
    if eng.the_prelude is None:
        eng.the_prelude = [(name, program(form)) for (name, _desc, form) in prelude.prelude]


class LiteBackend(object):
    """python/lib/shortcuts/__init__.py:113-117 (class Lite) without the front end"""
    def trace_constructor(self):
        from venture.lite.trace import Trace
        return Trace

    def name(self):
        return "lite"


def make_engine(backend=None, seed=1):
    """The reference Engine over `backend` (an object with trace_constructor() / name(), shortcuts/__init__.py:90-111)."""
    _import()
    _install_prelude_reader()
    from venture.engine.engine import Engine
    return Engine(backend if backend is not None else LiteBackend(), seed)


def expression(text):
    """model expression text -> desugared stack-dict expression (assume / observe / predict arguments)"""
    return desugar(read(text))


def value(x):
    """python scalar -> stack-dict value (observe)"""
    _import()
    import venture.value.dicts as val
    if isinstance(x, bool):
        return val.boolean(x)
    return val.number(x)
