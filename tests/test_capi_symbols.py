"""CPU: the C-ABI library loads and exports every symbol include/*.h declares (no compute calls without a GPU);
the ctypes mirror of fa_b200_params has the layout the C compiler gives the struct."""
import ctypes as C
import importlib
import os
import re
import subprocess
import sys
import tempfile

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
capi = importlib.import_module("xhy-flash-attention_b200._capi")


def _declared_symbols():
    names = []
    for hdr in ("fa_capi.h", "fa_b200.h"):
        src = open(os.path.join(ROOT, "include", hdr)).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        names += re.findall(r"^(?:bool|int|int64_t|void|const char \*)\s*\*?\s*(\w+)\s*\(", src, flags=re.M)
    return sorted(set(names))


@pytest.fixture(scope="module")
def lib_path():
    if not os.path.exists(capi.LIB_PATH):
        sys.path.insert(0, ROOT)
        import __graft_entry__ as g
        g.build()
    return capi.LIB_PATH


def test_exports_every_declared_symbol(lib_path):
    syms = _declared_symbols()
    assert len(syms) >= 20 and "flash_attn_fwd" in syms and "fa_b200_fwd" in syms
    out = subprocess.run(["nm", "-D", "--defined-only", lib_path], capture_output=True, text=True, check=True).stdout
    exported = {line.split()[-1] for line in out.splitlines() if line.strip()}
    missing = [s for s in syms if s not in exported]
    assert not missing, f"declared in include/*.h but not exported: {missing}"


def test_library_loads_and_reports_errors(lib_path):
    lib = C.CDLL(lib_path)
    lib.fa_b200_abi_version.restype = C.c_int
    assert lib.fa_b200_abi_version() == capi.FA_B200_ABI_VERSION
    lib.flash_attn_error.restype = C.c_char_p
    lib.flash_attn_set_error(b"boom")
    assert lib.flash_attn_error() == b"boom"
    lib.fa_b200_num_splits.restype = C.c_int
    assert lib.fa_b200_num_splits(64, 8, 4, 32768, 128, 148) == 1


def test_params_struct_layout_matches_c(lib_path):
    """Compile a tiny C program against include/fa_b200.h and compare sizeof/offsetof with the ctypes mirror."""
    fields = [f[0] for f in capi.Params._fields_]
    prog = ['#include <stdio.h>', '#include <stddef.h>', '#include "fa_b200.h"', 'int main(void){',
            'printf("%zu\\n", sizeof(fa_b200_params));']
    prog += [f'printf("%zu\\n", offsetof(fa_b200_params, {f}));' for f in fields]
    prog += ['return 0;}']
    with tempfile.TemporaryDirectory() as td:
        src = os.path.join(td, "t.c")
        open(src, "w").write("\n".join(prog))
        exe = os.path.join(td, "t")
        subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), "-I", "/usr/local/cuda/include", src, "-o", exe],
                       check=True)
        vals = [int(x) for x in subprocess.run([exe], capture_output=True, text=True, check=True).stdout.split()]
    assert vals[0] == C.sizeof(capi.Params)
    for f, off in zip(fields, vals[1:]):
        assert getattr(capi.Params, f).offset == off, f


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(capi, "_lib", None)
    monkeypatch.setattr(capi, "LIB_PATH", "/nonexistent/libflashattn_b200.so")
    with pytest.raises(RuntimeError, match="no CPU or PyTorch fallback"):
        capi.load()
