"""The C-ABI library loads and exports every symbol include/pbhc_b200.h declares (no compute calls)."""
import ctypes
import os
import re

from pbhc_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "pbhc_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(pbhc_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_header_symbols():
    names = _declared()
    assert len(names) >= 7
    L = ctypes.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(L, n), n
    assert sorted(_lib.EXPORTS) == names


def test_version_and_error_string():
    L = _lib.lib()
    assert L.pbhc_abi_version() == _lib.ABI_VERSION == 11
    assert isinstance(L.pbhc_last_error(), bytes)


def test_bad_arguments_return_codes():
    L = _lib.lib()
    assert L.pbhc_lookup(None, None, None, None, None, 0, None, None) == -1
    assert b"pbhc_lookup" in L.pbhc_last_error()
    assert L.pbhc_gae(None, None, None, None, 24, 8, 1, 0.99, 0.95, None, None, None, None) == -1
    assert L.pbhc_fk(None, None, None, 0, None, None, None, None) == -1
