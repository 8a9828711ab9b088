"""The C-ABI library loads and exports every symbol include/b200gpt.h declares (no compute calls: CPU-only box)."""
import ctypes
import os
import re

from lit_gpt_b200 import _cabi

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def header_functions():
    src = open(os.path.join(ROOT, "include", "b200gpt.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = _cabi.load()
    names = header_functions()
    assert len(names) >= 18
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/b200gpt.h but not exported by libb200gpt.so"
    # and the python binding covers exactly the header
    assert sorted(_cabi.exported_symbols()) == names


def test_abi_version_and_error_string():
    lib = _cabi.load()
    assert lib.b200_abi_version() == 1
    assert isinstance(lib.b200_last_error(), bytes)
    assert lib.b200_qgemv_workspace_bytes() > 0
    assert lib.b200_attn_decode_workspace_bytes(1, 32, 128, 8) > 32 * 8 * 128 * 4


def test_argument_validation_needs_no_gpu():
    """Bad arguments are rejected before any CUDA call, with a message."""
    lib = _cabi.load()
    rc = lib.b200_quantize_4bit(None, 0, 64, 64, 0, None, None, None)
    assert rc == 1 and b"null" in lib.b200_last_error()
    args = _cabi.QGemvArgs(batch=99)
    rc = lib.b200_qgemv(ctypes.byref(args), None)
    assert rc == 1
    assert lib.b200_qgemv(None, None) == 1


def test_no_torch_in_the_abi():
    """The boundary is plain C: the shared object must not link libtorch / libc10."""
    import subprocess

    out = subprocess.run(["ldd", _cabi.LIB_PATH], capture_output=True, text=True).stdout
    assert "torch" not in out and "c10" not in out
