"""The C-ABI library loads and exports every symbol include/velerode.h declares (no compute)."""
import ctypes
import os
import re

from util import ROOT


def _declared():
    txt = open(os.path.join(ROOT, "include", "velerode.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(ve_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    path = os.path.join(ROOT, "veloren_b200", "csrc", "libvelerode.so")
    if not os.path.exists(path):
        from veloren_b200 import erosion
        erosion.build()
    lib = ctypes.CDLL(path)
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), "missing export %s" % n


def test_python_mirror_lists_the_same_exports():
    from veloren_b200 import erosion
    assert sorted(erosion.EXPORTS) == [n for n in _declared() if n in erosion.EXPORTS]
    missing = [n for n in _declared() if n not in erosion.EXPORTS]
    assert missing == [], missing


def test_no_cpu_fallback_in_product():
    """The product package must not reference the oracle."""
    pkg = os.path.join(ROOT, "veloren_b200")
    for d, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(d, fn), errors="ignore").read()
                assert "pyoracle" not in src and "erosion_oracle" not in src, fn
