"""The parallelisable formulation of the lake growth (tools/lake_growth_prototype.py: Kruskal
reconstruction tree + direction-aware choice inside a weight class) must reproduce the reference's
serial heap loop (erosion.rs:2066-2137, as restated by the oracle): same pass per lake, same order."""
import os
import sys
import tempfile

import numpy as np
import pytest

from util import ROOT
sys.path.insert(0, os.path.join(ROOT, "tools"))
import lake_growth_prototype as lg          # noqa: E402
from veloren_b200 import synth              # noqa: E402
from oracle import pyoracle as po           # noqa: E402


@pytest.mark.parametrize("kind", [0, 1])
def test_krt_growth_equals_serial_prim(kind, monkeypatch):
    sys.setrecursionlimit(200000)
    f = synth.generate(7, 7, kind=kind)
    o = po.Oracle(f)
    path = os.path.join(tempfile.gettempdir(), "orc_lakes_dump_test_%d_%d.bin" % (os.getpid(), kind))
    total = 0
    for it in range(2):
        h_before = o.get("h").copy()
        monkeypatch.setenv("ORC_DUMP_LAKES", path)
        o.step()
        monkeypatch.delenv("ORC_DUMP_LAKES")
        edges, conn = lg.read_dump(path)
        ref = [tuple(int(v) for v in r) for r in conn]
        assert lg.grow(edges, h_before) == ref, "step %d" % it
        total += len(ref)
    os.remove(path)
    assert total > 50          # the case has lakes at all
