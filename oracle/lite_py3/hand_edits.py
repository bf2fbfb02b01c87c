"""Per-file one-line py2 -> py3 edits the regex pass of build_lite_py3.py cannot express.
Each entry is (anchor text, replacement); the anchor is the shortest fragment that locates the
site.  All are syntax or iterator-laziness fixes; none changes what the reference computes."""

HAND = {
    # three-argument raise spread over two lines
    "lite/inference_sps.py": [
        ("err, info, self.addr), None, info[2]", "err, info, self.addr).with_traceback(info[2])"),
    ],
    "lite/regen.py": [
        ("address=address), \\\n        None, info[2]", "address=address).with_traceback(info[2])"),
        ("cause=err), None, info[2]", "cause=err).with_traceback(info[2])"),
    ],
    # py2 int / int is floor division (scaffolder_loop, dispatch.py:127; ct is a sum of ints)
    "lite/infer/dispatch.py": [
        ("return ct/len(scaffolders)", "return ct//len(scaffolders)"),
    ],
    # print statement on a def line
    "engine/inference.py": [
        ("def printf(self, dataset): print dataset.asPandas()", "def printf(self, dataset): print(dataset.asPandas())"),
    ],
    # py3 laziness: map() results that are concatenated to lists / pattern-matched as lists
    "sivm/macro.py": [
        ("pattern = ['cond'] + map(list, zip(preds, exprs))", "pattern = ['cond'] + list(map(list, zip(preds, exprs)))"),
        ("pattern = ['let', map(list, zip(syms, vals)), 'body']", "pattern = ['let', list(map(list, zip(syms, vals))), 'body']"),
        ("pattern = ['letrec', map(list, zip(syms, vals)), 'body']", "pattern = ['letrec', list(map(list, zip(syms, vals))), 'body']"),
    ],
    # py2 ordered objects through __cmp__; py3's sorted() asks for __lt__ (getOrderedSetsInScope sorts block keys,
    # lite/trace.py:334): the same total order, spelled the py3 way
    "lite/value.py": [
        ("  def __cmp__(self, other):", "  def __lt__(self, other):\n    return self.__cmp__(other) < 0\n  def __cmp__(self, other):"),
    ],
    "sivm/core_sivm.py": [
        ("return map(_modify_expression, expression)", "return list(map(_modify_expression, expression))"),
    ],
}
