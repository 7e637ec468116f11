"""CPU: the C-ABI library builds, loads and exports every symbol include/lg_b200.h declares
(no compute calls without a GPU)."""
import ctypes
import os
import re

from conftest import ROOT, pkg


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "lg_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(lg_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    C = pkg("_cabi")
    if not os.path.isfile(C.LIB_PATH):
        C.build()
    lib = ctypes.CDLL(C.LIB_PATH)
    declared = _declared_symbols()
    assert declared, "no symbols parsed from the header"
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/lg_b200.h but not exported"
    # and the binding table covers the header exactly
    assert sorted(C.SIGNATURES) == declared


def test_version_and_pure_host_entry_points():
    C = pkg("_cabi")
    assert C.version() == 100
    assert C.lib().lg_filter_ws_bytes(2, 10, 20) == 2 * 30 * 8
    assert C.tiles_pad(1) == 2 and C.tiles_pad(256) == 2 and C.tiles_pad(257) == 4
