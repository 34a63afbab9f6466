"""The ctypes structures of ops/_native.py must have the layout the C header declares: a silent mismatch
would hand kernels garbage pointers.  Compiles a tiny C++ program against include/fps_b200.h (host
compiler only, no GPU) and compares sizeof / offsetof with ctypes."""
import os
import shutil
import subprocess

import pytest

from fps_b200.ops import _native as N

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

STRUCTS = {
    "FpsTableView": N.FpsTableView,
    "FpsFlow": N.FpsFlow,
    "FpsMfParams": N.FpsMfParams,
    "FpsPaParams": N.FpsPaParams,
    "FpsNegSampler": N.FpsNegSampler,
}
RENAMED = {"lambda_": "lambda"}          # Python keyword


@pytest.mark.skipif(shutil.which("g++") is None, reason="needs a host C++ compiler")
def test_ctypes_structs_match_the_c_header(tmp_path):
    lines = ['#include <cstdio>', '#include <cstddef>', f'#include "{ROOT}/include/fps_b200.h"', "int main() {"]
    for name, cls in STRUCTS.items():
        lines.append(f'  printf("{name} sizeof %zu\\n", sizeof({name}));')
        for field, _ in cls._fields_:
            c_field = RENAMED.get(field, field)
            lines.append(f'  printf("{name} {field} %zu\\n", offsetof({name}, {c_field}));')
    lines += ["  return 0;", "}"]
    src = tmp_path / "layout.cpp"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run(["g++", "-std=c++17", str(src), "-o", str(exe)], check=True, capture_output=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout
    got = {}
    for line in out.splitlines():
        s, f, v = line.split()
        got[(s, f)] = int(v)
    for name, cls in STRUCTS.items():
        import ctypes as C
        assert got[(name, "sizeof")] == C.sizeof(cls), name
        for field, _ in cls._fields_:
            assert got[(name, field)] == getattr(cls, field).offset, (name, field)
    assert N.FLOW_BYTES == got[("FpsFlow", "sizeof")]


def _header_prototypes():
    import re
    text = open(os.path.join(ROOT, "include", "fps_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    protos = {}
    for m in re.finditer(r"\b(?:int|size_t|const char\*)\s+(fps_\w+)\s*\(([^;{}]*)\)\s*;", text):
        args = m.group(2).strip()
        protos[m.group(1)] = 0 if args in ("", "void") else args.count(",") + 1
    return protos


def test_every_declared_entry_point_is_exported_with_the_bound_arity():
    """Header prototype <-> exported symbol <-> ctypes argtypes (loads the library; needs no GPU)."""
    if not N.available():
        pytest.skip("libfps_b200.so has not been built")
    try:
        L = N.lib()
    except OSError as e:                      # e.g. no CUDA runtime on the machine
        pytest.skip(f"cannot load the library here: {e}")
    protos = _header_prototypes()
    assert len(protos) >= 30
    for name, n_args in protos.items():
        fn = getattr(L, name, None)
        assert fn is not None, f"{name} is declared in fps_b200.h but not exported"
        assert fn.argtypes is not None, f"{name} is declared in fps_b200.h but has no ctypes signature in ops/_native.py"
        assert len(fn.argtypes) == n_args, (name, len(fn.argtypes), n_args)
