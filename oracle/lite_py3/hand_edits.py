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
    # print statement on a def line
    "engine/inference.py": [
        ("def printf(self, dataset): print dataset.asPandas()", "def printf(self, dataset): print(dataset.asPandas())"),
    ],
}
