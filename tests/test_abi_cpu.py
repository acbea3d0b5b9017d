"""CPU: the C-ABI library loads and exports every symbol include/mpcb200.h declares (no compute calls)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "mpcb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mpcb_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported_and_bound():
    import mpcb200
    from mpcb200 import _ffi
    names = _declared_symbols()
    assert len(names) >= 18
    L = ctypes.CDLL(_ffi.LIB_PATH)
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/mpcb200.h but not exported"
        assert n in _ffi.SIGNATURES, f"{n} has no ctypes signature in _ffi.py"
    assert _ffi.lib().mpcb_version() == 100


def test_struct_layouts_match_header_defaults():
    """ctypes mirrors of MpcProblemDesc / MpcSolverParams agree with the C side field by field."""
    from mpcb200 import _ffi
    from oracle import oracle as orc
    d, d2 = _ffi.default_desc(0, 20, 100), orc.default_desc(0, 20, 100)
    for name, _ in _ffi.MpcProblemDesc._fields_:
        a, b = getattr(d, name), getattr(d2, name)
        assert (list(a) == list(b)) if hasattr(a, "__len__") else (a == b), name
    s, s2 = _ffi.default_solver_params(20), orc.default_solver_params(20)
    for name, _ in _ffi.MpcSolverParams._fields_:
        assert getattr(s, name) == getattr(s2, name), name
    assert (_ffi.default_vehicle_params() == orc.DEFAULT_P).all()


def test_no_cpu_fallback():
    """Without a CUDA device the product path raises; it never computes on the host."""
    import torch
    import mpcb200
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(mpcb200.MpcbError):
        mpcb200.BatchedMpc(N=12)
    from mpcb200 import _ffi
    h = ctypes.c_void_p()
    d = _ffi.default_desc()
    assert _ffi.lib().mpcb_create(ctypes.byref(d), ctypes.byref(h)) == _ffi.E_NODEVICE


def test_product_never_imports_oracle():
    """The product path must not import, include, link or call anything under oracle/."""
    pats = [r"^\s*(from|import)\s+oracle", r"#include\s*[<\"][^>\"]*oracle", r"libmpc_oracle", r"\borc_[a-z_]+\s*\(",
            r"namespace\s+orc"]
    pkg = os.path.join(ROOT, "model-predictive-control_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                for pat in pats:
                    assert not re.search(pat, src, flags=re.M), f"{f} matches {pat}: product must not use the oracle"


def test_dedupe_roads_host_logic():
    """alpaqa_compat.dedupe_roads: per-instance centerlines that take a few distinct values become a road table + index
    (the shared-memory pruned scan), anything else stays per-instance (None)."""
    import numpy as np
    from mpcb200 import alpaqa_compat as pa
    rng = np.random.default_rng(0)
    t = rng.normal(size=(3, 40))
    idx = rng.integers(0, 3, 500)
    table, got = pa.dedupe_roads(t[idx])
    assert table.shape[0] == 3 and got.dtype == np.int32
    assert np.array_equal(table[got], t[idx])                       # every row maps to its own road
    assert np.array_equal(table[0], t[idx[0]])                      # roads are numbered by first appearance
    one, i1 = pa.dedupe_roads(np.tile(t[1], (7, 1)))
    assert one.shape == (1, 40) and (i1 == 0).all()
    assert pa.dedupe_roads(rng.normal(size=(pa.MAX_TABLE_ROADS + 1, 40))) is None   # too many distinct roads
    bad = np.tile(t[0], (4, 1))
    bad[2, 5] = np.nan                                              # a NaN row equals nothing, not even itself
    assert pa.dedupe_roads(np.tile(bad, (pa.MAX_TABLE_ROADS + 1, 1))) is None
