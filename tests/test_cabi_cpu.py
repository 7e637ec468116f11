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


def test_library_is_built_from_this_tree(tmp_path, monkeypatch):
    """lg_source_id(): the digest the Makefile embeds equals the digest of csrc/ + the header in the tree, and a
    library with another digest is not loaded silently (the binding rebuilds it or raises)."""
    C = pkg("_cabi")
    assert C.lib().lg_source_id().decode() == C.source_id() == C._built_id()
    stale = tmp_path / "liblg_b200.so"
    data = open(C.LIB_PATH, "rb").read().replace(b"LG_SOURCE_ID=" + C.source_id().encode(), b"LG_SOURCE_ID=" + b"0" * 16)
    stale.write_bytes(data)
    monkeypatch.setattr(C, "LIB_PATH", str(stale))
    monkeypatch.setattr(C, "_lib", None)
    monkeypatch.setattr(C, "build", lambda verbose=False: None)  # "rebuilding" leaves the stale file in place
    monkeypatch.delenv("LG_B200_LIB", raising=False)
    import pytest
    with pytest.raises(RuntimeError, match="stale"):
        C.lib()
