"""CPU: the C-ABI library loads and exports every symbol include/az_b200.h declares; the ctypes mirror of
`struct az_tree` matches the C layout.  No compute calls (there is no GPU here)."""
import ctypes
import os
import re
import subprocess

from alphazero_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "az_b200.h")


def _declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(az_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    if _lib.library_is_stale():
        _lib.build_library()
    L = ctypes.CDLL(_lib.LIB_PATH)
    names = _declared_symbols()
    assert len(names) >= 16
    for name in names:
        assert hasattr(L, name), f"{name} declared in az_b200.h but not exported"
        assert name in _lib.SIGNATURES, f"{name} has no ctypes signature in alphazero_b200/_lib.py"
    assert set(_lib.SIGNATURES) == set(names)
    assert _lib.lib().az_abi_version() == 10


def test_struct_layout_matches_c(tmp_path):
    prog = tmp_path / "layout.c"
    fields = [f[0] for f in _lib.AzTree._fields_]
    body = "".join(f'printf("%zu\\n", offsetof(az_tree, {f}));' for f in fields)
    prog.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "az_b200.h"\n'
                    f'int main(void){{printf("%zu\\n", sizeof(az_tree));printf("%zu\\n", sizeof(az_node));{body}return 0;}}')
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(prog)], check=True)
    out = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    assert out[0] == ctypes.sizeof(_lib.AzTree)
    assert out[1] == 32
    for f, off in zip(fields, out[2:]):
        assert getattr(_lib.AzTree, f).offset == off, f


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libaz_b200.so")
    try:
        _lib.lib()
    except RuntimeError as e:
        assert "no CPU fallback" in str(e)
    else:
        raise AssertionError("loading a missing library must raise")
