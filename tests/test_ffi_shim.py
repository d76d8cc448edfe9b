"""The XLA FFI shim (ffi/jm_ffi.cc) cannot be compiled in this image (no jaxlib headers); what can be checked on CPU:
the file pre-processes and compiles as C++ in its gated form, every handler names a function the C ABI header
declares, and the Python registration list matches the handler symbols."""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "ffi", "jm_ffi.cc")


def test_shim_compiles_in_its_gated_form(tmp_path):
    obj = tmp_path / "jm_ffi.o"
    res = subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-I", os.path.join(ROOT, "include"), SRC],
                         capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    res = subprocess.run(["g++", "-std=c++17", "-c", "-fPIC", "-I", os.path.join(ROOT, "include"), SRC, "-o", str(obj)],
                         capture_output=True, text=True)
    assert res.returncode == 0 and obj.exists(), res.stderr


def test_handlers_call_declared_entry_points_and_match_the_python_list():
    src = open(SRC).read()
    header = open(os.path.join(ROOT, "include", "jumbo_mae_b200.h")).read()
    declared = set(re.findall(r"\b(jm_[a-z0-9_]+)\s*\(", header))
    called = set(re.findall(r"\b(jm_[a-z0-9_]+)\s*\(", src)) - {"jm_ffi_unavailable"}
    assert called and called <= declared, called - declared
    symbols = re.findall(r"XLA_FFI_DEFINE_HANDLER_SYMBOL\((\w+),", src)
    assert len(symbols) == len(set(symbols)) >= 20
    ns = {}
    exec(compile(open(os.path.join(ROOT, "ffi", "jm_ffi.py")).read().split("def register")[0], "jm_ffi.py", "exec"), ns)
    assert ns["TARGETS"] == symbols
    # every handler body stays a thin shim
    bodies = re.findall(r"static ffi::Error \w+\([^)]*\)\s*\{(.*?)\n\}", src, flags=re.S)
    assert bodies and max(b.count("\n") for b in bodies) <= 20
