"""CPU checks of the C-ABI boundary: the library builds for sm_100a, loads, and exports
every symbol include/jumbo_mae_b200.h declares.  No compute call is made here."""
import os
import re

from jumbo_mae_tpu_b200 import _lib, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "jumbo_mae_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(jm_[a-z0-9_]+)\s*\(", text)))


def test_library_builds_and_exports_every_declared_symbol():
    build.build()
    lib = _lib.load()
    declared = _declared_symbols()
    assert declared, "no declarations parsed from the header"
    missing = [s for s in declared if not hasattr(lib, s)]
    assert not missing, f"library does not export: {missing}"
    assert sorted(_lib.EXPORTS) == declared
    assert lib.jm_version() >= 100


def test_errors_are_reported_not_thrown():
    lib = _lib.load()
    # null args -> non-zero return code and a message, no crash, no GPU needed
    rc = lib.jm_gemm(None, None)
    assert rc != 0
    assert b"jm_gemm" in lib.jm_last_error()
    rc = lib.jm_mask_sort(None, 5000, 10, None, None, None, None)
    assert rc != 0 and b"length" in lib.jm_last_error()
