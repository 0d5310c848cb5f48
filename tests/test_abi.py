"""The C-ABI library loads and exports every symbol include/universe_b200.h declares
(no compute calls: there is no GPU on the CPU tier)."""
import ctypes
import os
import re

from conftest import ROOT
from open_universe_b200 import _abi, _native


def _declared_functions():
    src = open(os.path.join(ROOT, "include", "universe_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(uv_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = _native.lib()
    names = _declared_functions()
    assert len(names) >= 13
    for n in names:
        assert hasattr(lib, n), n
    assert sorted(_native.EXPORTS) == names


def test_abi_version_and_record_sizes():
    lib = _native.lib()
    assert lib.uv_abi_version() == _abi.UV_ABI_VERSION
    assert lib.uv_sizeof_op() == ctypes.sizeof(_abi.UvOp)
    assert ctypes.sizeof(_abi.UvStepCoef) == 16 and ctypes.sizeof(_abi.UvTens) == 32


def test_invalid_arguments_return_error_codes():
    lib = _native.lib()
    h = ctypes.c_void_p()
    assert lib.uv_model_create(ctypes.byref(h), None, 0, 0) == -1
    assert b"null" in lib.uv_last_error()
    assert lib.uv_enhance(None, None, None, None, None, None, 0, None) == -1
