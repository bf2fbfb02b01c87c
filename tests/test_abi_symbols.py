"""The C-ABI library loads without a GPU and exports every symbol that
include/venture_b200.h declares; compute calls fail loudly without a device
(no CPU fallback)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "venture_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(vk_[a-z0-9_]+)\s*\(", src)))


def test_library_builds_and_exports_all_declared_symbols():
    from venturecxx_b200 import build, _lib
    build.build()
    lib = ctypes.CDLL(_lib.SO_PATH)
    names = declared_symbols()
    assert len(names) > 30
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    # the ctypes binding covers the same surface
    assert sorted(_lib.PROTOTYPES) == names
    assert _lib.load().vk_version() >= 100


def test_no_cpu_fallback_without_device():
    import venturecxx_b200 as vb
    from venturecxx_b200 import _lib
    n = ctypes.c_int(-1)
    _lib.call("vk_device_count", ctypes.byref(n))
    if n.value > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(vb.VentureException) as e:
        vb.B200TraceSet(vb.NormalNormalModel(), 8)
    assert "no CPU fallback" in str(e.value)


def test_product_package_never_imports_oracle():
    pkg = os.path.join(ROOT, "venturecxx_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert "import oracle" not in txt and "liboracle" not in txt, f
