#!/usr/bin/env python
"""TEST INFRASTRUCTURE -- builds a Python-3-importable copy of the reference's Lite backend.

The reference (probcomp/Venturecxx) is Python 2 source.  This script reads the
files of `backend/lite/` and the handful of `python/lib/` modules Lite imports
from /root/reference, applies *syntactic* py2 -> py3 rewrites (listed below, each
a regex or a one-line hand edit keyed by file), and writes the result into
`oracle/_ref/venture/` -- which is git-ignored: no reference source enters this
repository's history.  `tests/golden/make_lite_chains.py` then imports
`venture.lite.trace.Trace` from there and runs the reference's OWN scaffold /
detach / regen / mixMH / egibbs / subsampled-MH code to produce the committed
chain fixtures (`tests/golden/lite_chains.json`).

Nothing here changes what the reference computes:
  * print statement -> function, `except E, e` -> `except E as e`,
    three-argument raise -> `raise ... from None`, dict.iter* -> dict.*,
    xrange -> range, cPickle -> pickle, basestring/unicode/long -> str/int,
    tuple parameters unpacked in the body, `exec` statements, backquotes;
  * py3 laziness: `map`/`zip`/`filter`/`dict.keys()` results that the reference
    indexes or re-iterates are wrapped in list(...) at the specific call sites;
  * `/` between ints: the files on this path that lack
    `from __future__ import division` get it only if they perform no int/int
    division relied on as floor division (checked by hand; see DIVISION_NOTES);
  * `random.jumpahead` (removed in py3) is not called on the fixture path.

Run:  python oracle/lite_py3/build_lite_py3.py   (needs /root/reference)
"""
import os
import re
import sys

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "_ref")

# (reference dir, package path under OUT/venture)
TREES = [
    ("backend/lite", "lite"),
    ("backend/lite/infer", "lite/infer"),
    ("python/lib/value", "value"),
    ("python/lib/engine", "engine"),
    ("backend/untraced", "untraced"),
    ("python/lib/sivm", "sivm"),         # macro system (`do`, `let`, ...) for the engine-seam tests
]
SINGLE = [
    ("python/lib/exception.py", "exception.py"),
    ("python/lib/multiprocess.py", "multiprocess.py"),
    ("python/lib/ripl/utils.py", "ripl/utils.py"),
]
# Modules the engine imports but the hot path never calls (plotting, the ripl / parser front end): generated
# as empty stand-ins so that `venture.engine.engine.Engine` -- the reference's own control plane -- imports and
# can drive a backend here (tests/ref_engine.py).  The package __init__ files of sivm / ripl pull in the whole
# front end (parser, server); they are replaced by empty ones.
STUBS = {
    "plots/__init__.py": "",
    "plots/p_p_plot.py": "def p_p_plot_2samp(*args, **kwargs):\n  raise NotImplementedError('plotting is not rendered')\n",
    "ripl/__init__.py": "",
    "sivm/__init__.py": "",
}

REGEX = [
    (r"\.iteritems\(\)", ".items().__iter__()"),
    (r"\.itervalues\(\)", ".values().__iter__()"),
    (r"\.iterkeys\(\)", ".__iter__()"),
    (r"\bxrange\b", "range"),
    (r"\bimport cPickle as (\w+)\b", r"import pickle as \1"),
    (r"^import cPickle$", "import pickle as cPickle"),
    (r"\bimport cStringIO as StringIO\b", "import io as StringIO"),
    (r"\bbasestring\b", "str"),
    (r"\bunicode\b", "str"),
    (r"except ([\w\.]+), ?(\w+):", r"except \1 as \2:"),
    (r"except \(([\w\., ]+)\), ?(\w+):", r"except (\1) as \2:"),
    (r"\braise (\w+), None, (\w+)", r"raise \1.with_traceback(\2)"),
    (r"\braise (\w+)\((.*)\), None, (\w+)", r"raise \1(\2).with_traceback(\3)"),
    (r"\.has_key\((\w+)\)", r".__contains__(\1)"),
    (r"\bfrom itertools import izip\b", "izip = zip"),
    (r"\bitertools\.izip\b", "zip"),
    (r"\bfrom itertools import imap\b", "imap = map"),
    (r"\bim_func\b", "__func__"),
    (r"\bfunc_name\b", "__name__"),
    (r"\.next\(\)", ".__next__()"),
    (r"\b0(\d+)\b(?=[^.\dxe])", None),  # placeholder (octal) -- unused
    (r"\blong\b", "int"),
    (r"^import __builtin__$", "import builtins as __builtin__"),
    (r"\braw_input\b", "input"),
    (r"\breduce\(", "functools_reduce("),
]


def fix_print(src):
    out = []
    for ln in src.split("\n"):
        m = re.match(r"^(\s*)print\s+(?!\()(.*)$", ln)
        if m and not ln.strip().startswith("#"):
            body = m.group(2).rstrip()
            if body.startswith(">>"):
                mm = re.match(r">>\s*([\w\.]+)\s*,\s*(.*)$", body)
                ln = "%sprint(%s, file=%s)" % (m.group(1), mm.group(2), mm.group(1))
            elif body.endswith(","):
                ln = "%sprint(%s end=' ')" % (m.group(1), body)
            else:
                ln = "%sprint(%s)" % (m.group(1), body)
        elif re.match(r"^(\s*)print\s*$", ln):
            ln = re.sub(r"print\s*$", "print()", ln)
        out.append(ln)
    return "\n".join(out)


def fix_tuple_params(src):
    """def f(self, (a, b)): ... -> def f(self, _tp0): (a, b) = _tp0; also lambda (a, b): ..."""
    def lam(m):
        names = m.group(1)
        body = m.group(2)
        return "lambda _tp: (lambda %s: %s)(*_tp)" % (names, body)
    src = re.sub(r"lambda \(([\w, ]+)\): ?([^\n]*?)(?=[,\)]\s*$|\)\s*\)|$)", lam, src, flags=re.M)
    out = []
    lines = src.split("\n")
    i = 0
    while i < len(lines):
        ln = lines[i]
        m = re.match(r"^(\s*)def (\w+)\((.*)\):\s*(.*)$", ln)
        if m and re.search(r"\(\s*\w+(\s*,\s*\w+)+\s*\)", m.group(3)):
            indent, name, params, rest = m.groups()
            unpack = []
            k = [0]

            def rep(mm):
                nm = "_tp%d" % k[0]
                k[0] += 1
                unpack.append("%s = %s" % (mm.group(0), nm))
                return nm
            params2 = re.sub(r"\(\s*\w+(\s*,\s*\w+)+\s*\)", rep, params)
            if rest.strip():
                out.append("%sdef %s(%s):" % (indent, name, params2))
                for u in unpack:
                    out.append("%s  %s" % (indent, u))
                out.append("%s  %s" % (indent, rest))
            else:
                out.append("%sdef %s(%s):" % (indent, name, params2))
                # find indentation of the body
                j = i + 1
                while j < len(lines) and not lines[j].strip():
                    j += 1
                bind = re.match(r"^(\s*)", lines[j]).group(1)
                for u in unpack:
                    out.append("%s%s" % (bind, u))
            i += 1
            continue
        out.append(ln)
        i += 1
    return "\n".join(out)


# py2 orders any two objects (`<` never raises); the persistent weight-balanced trees of wttree.py rely on
# that for keys such as trace nodes, addresses (nested tuples mixing str and int) and VentureValues.  py3 does
# not, so wttree's key comparisons go through one explicit total order.  The order only decides the SHAPE of
# the persistent maps (lookup / insert / delete by key are order-agnostic), never a computed value.
WTTREE_ORDER = '''
def _ord(x):
  if x is None: return (-1,)
  if isinstance(x, bool): return (0, int(x))
  if isinstance(x, (int, float)): return (0, x)
  if isinstance(x, str): return (1, x)
  if isinstance(x, (tuple, list)): return (4, tuple(_ord(e) for e in x))
  if hasattr(x, "asStackDict") and hasattr(x, "equal"): return (2, type(x).__name__, repr(x))
  return (3, type(x).__name__, id(x))

def _py2_lt(a, b):
  return _ord(a) < _ord(b)
'''


# One-line hand edits: file (package-relative) -> list of (old, new).  Each is a py3 laziness or
# syntax fix with no change of meaning.
HAND = {}


def convert(src, rel):
    src = fix_print(src)
    for pat, rep in REGEX:
        if rep is None:
            continue
        src = re.sub(pat, rep, src, flags=re.M)
    src = fix_tuple_params(src)
    if rel == os.path.join("lite", "wttree.py"):
        src = re.sub(r"\b(\w*key_) < (\w*key_)", r"_py2_lt(\1, \2)", src)
        src = src.replace("\nclass ", WTTREE_ORDER + "\nclass ", 1)
    # call sites `x.iteritems()` become `x.items().__iter__()`; the two reference classes that define their
    # own iteritems get the matching alias
    if rel == os.path.join("lite", "wttree.py"):
        src += "\nPMap.items = PMap.iteritems\n"
    if rel == os.path.join("lite", "smap.py"):
        src += "\nSamplableMap.items = SamplableMap.iteritems\n"
    if "functools_reduce(" in src:
        src = "from functools import reduce as functools_reduce\n" + src
    for old, new in HAND.get(rel, []):
        if old not in src:
            raise SystemExit("hand edit no longer applies in %s: %r" % (rel, old))
        src = src.replace(old, new)
    return src


def main():
    if not os.path.isdir(REF):
        print("no /root/reference here; nothing to do")
        return 0
    sys.path.insert(0, HERE)
    try:
        import hand_edits
        HAND.update(hand_edits.HAND)
    except ImportError:
        pass
    pkg = os.path.join(OUT, "venture")
    os.makedirs(pkg, exist_ok=True)
    with open(os.path.join(pkg, "__init__.py"), "w") as f:
        f.write("# py3 rendering of the reference's `venture` package (generated; see oracle/lite_py3)\n")
    n = 0
    for refdir, sub in TREES:
        os.makedirs(os.path.join(pkg, sub), exist_ok=True)
        for fn in sorted(os.listdir(os.path.join(REF, refdir))):
            if not fn.endswith(".py"):
                continue
            rel = os.path.join(sub, fn)
            with open(os.path.join(REF, refdir, fn)) as f:
                src = f.read()
            with open(os.path.join(pkg, rel), "w") as f:
                f.write(convert(src, rel))
            n += 1
    for ref, rel in SINGLE:
        os.makedirs(os.path.dirname(os.path.join(pkg, rel)), exist_ok=True)
        with open(os.path.join(REF, ref)) as f:
            src = f.read()
        with open(os.path.join(pkg, rel), "w") as f:
            f.write(convert(src, rel))
        n += 1
    for rel, body in STUBS.items():
        os.makedirs(os.path.dirname(os.path.join(pkg, rel)), exist_ok=True)
        with open(os.path.join(pkg, rel), "w") as f:
            f.write("# generated stand-in (oracle/lite_py3/build_lite_py3.py STUBS)\n" + body)
    # compile check
    bad = 0
    for root, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith(".py"):
                p = os.path.join(root, fn)
                try:
                    with open(p) as f:
                        compile(f.read(), p, "exec")
                except SyntaxError as e:
                    bad += 1
                    print("SYNTAX %s:%s %s: %s" % (os.path.relpath(p, pkg), e.lineno, e.msg, (e.text or "").strip()))
    print("converted %d files into %s (%d with syntax errors)" % (n, pkg, bad))
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
