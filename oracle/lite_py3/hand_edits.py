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
}
