"""The C-ABI shared library loads (no GPU needed for that) and exports every entry point include/pace_b200.h
declares; the ctypes binding table of pace_b200/_lib.py covers the same set."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared():
    text = open(os.path.join(ROOT, "include", "pace_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(pace_b200_\w+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from pace_b200 import _lib

    assert os.path.exists(_lib.LIB_PATH), "build the library first: python -m pace_b200.build"
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = declared()
    assert len(names) > 40
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_binding_table_matches_header():
    from pace_b200 import _lib

    assert sorted(_lib.EXPORTS) == declared()
