"""The C-ABI library loads on a CPU-only box and exports every symbol that
include/drtvam_b200.h declares (no compute calls here)."""
import ctypes as C
import os
import re

from drtvam_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    with open(os.path.join(ROOT, "include", "drtvam_b200.h")) as f:
        src = f.read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(drtvam_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    names = declared_symbols()
    assert len(names) >= 14
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"
        assert n in _lib.SYMBOLS, f"{n} has no ctypes prototype"


def test_struct_layout_matches_header():
    # sizeof is validated by the library at scene creation; here the static parts
    assert C.sizeof(_lib.DrtvamSurface) == 48
    assert _lib.DrtvamSceneDesc.surf.size == 48 * _lib.MAX_SURFACES
    assert _lib.load().drtvam_abi_version() == _lib.ABI_VERSION


def test_no_oracle_or_cpu_fallback_in_product():
    """The product package must never import the oracle."""
    pkg = os.path.join(ROOT, "drtvam_b200")
    for base, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h")):
                with open(os.path.join(base, fn)) as f:
                    text = f.read()
                assert "import oracle" not in text and "liboracle" not in text, fn
