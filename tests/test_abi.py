"""The C-ABI library on a machine without a GPU: it builds, loads, exports every symbol
include/ctb_clib.h declares, follows the reference CLib's error conventions, and refuses to run
without a CUDA device (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import MECHS, ROOT


@pytest.fixture(scope="module")
def lib():
    from cantera_b200 import build, clib
    build.build()
    return clib.lib()


def _declared():
    hdr = open(os.path.join(ROOT, "include", "ctb_clib.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(ctb_[A-Za-z0-9_]+)\s*\(", hdr)))


def test_header_symbols_exported(lib):
    from cantera_b200 import clib
    names = _declared()
    assert len(names) >= 35
    for n in names:
        assert hasattr(lib, n), n
    assert sorted(clib.SIGNATURES) == names


def test_bridge_exports(lib):
    """libctb_cantera.so (include/ctb_cantera.h), where it was built: every declared symbol is exported,
    and a bad file name comes back as -1 with the reference's message on the error stack."""
    from cantera_b200 import clib
    b = clib.bridge()
    if b is None:
        pytest.skip("libctb_cantera.so not built (needs the reference)")
    hdr = open(os.path.join(ROOT, "include", "ctb_cantera.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = sorted(set(re.findall(r"\b(ctb_[A-Za-z0-9_]+)\s*\(", hdr)))
    assert names == sorted(clib.BRIDGE_SIGNATURES)
    for n in names:
        assert hasattr(b, n), n
    assert b.ctb_kin_new(-1, b"/nonexistent-mechanism.yaml", b"", 0) == -1
    assert "nonexistent-mechanism" in clib.last_error()
    assert b.ctb_kin_new(3, b"gri30.yaml", b"", 0) == -1


def test_version_and_string_convention(lib):
    need = lib.ctb_version(0, None)  # (0, NULL) -> required length (test/clib/test_ctkin.cpp:119-123)
    buf = C.create_string_buffer(need)
    assert lib.ctb_version(need, buf) == need
    assert b"sm_100a" in buf.value


def test_bad_handle_and_error_stack(lib):
    from cantera_b200 import clib
    assert lib.ctb_kin_nSpecies(12345) == -1
    assert "out of range" in clib.last_error()
    assert lib.ctb_kin_multiplier(12345, 0) == clib.DERR
    assert lib.ctb_kin_del(12345) == -1
    assert lib.ctb_reactornet_advance(777, 1.0) == -1
    assert lib.ctb_kin_newFromFile(b"/nonexistent.ctbm", 0) == -1
    assert "cannot open" in clib.last_error()


def test_no_cpu_fallback(lib):
    """Without a usable CUDA device every constructor fails loudly."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from cantera_b200 import BatchedGasKinetics, CtbError
    with pytest.raises(CtbError, match="no usable CUDA device|CUDA"):
        BatchedGasKinetics("gri30")


def test_product_does_not_touch_oracle():
    """Nothing under cantera_b200/ or include/ may import, include or link the oracle."""
    bad = []
    for root in ("cantera_b200", "include"):
        for d, _, files in os.walk(os.path.join(ROOT, root)):
            if "build" in d.split(os.sep):
                continue
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                    txt = open(os.path.join(d, f), errors="replace").read()
                    if re.search(r"ctb_oracle|oracle_py|libctb_oracle|from oracle|import oracle", txt):
                        bad.append(os.path.join(d, f))
    assert not bad, bad


def test_ctbm_roundtrip(tmp_path):
    from cantera_b200.ctbm import read_arrays, write_arrays
    a = read_arrays(MECHS["h2o2"])
    assert a["dims"].tolist() == [10, 29]
    p = tmp_path / "x.ctbm"
    write_arrays(str(p), a)
    b = read_arrays(str(p))
    assert sorted(a) == sorted(b)
    for k in a:
        if isinstance(a[k], str):
            assert a[k] == b[k]
        else:
            assert np.array_equal(a[k], b[k])
