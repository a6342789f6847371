"""The C-ABI library loads (without a GPU) and exports every symbol include/b200ep.h declares."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    src = open(os.path.join(ROOT, "include", "b200ep.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200ep_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from b200ep import _lib

    lib = _lib.load()
    syms = _header_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/b200ep.h but not exported"
    assert sorted(_lib.SIGNATURES) == syms, "ctypes SIGNATURES and the header disagree"
    assert lib.b200ep_version() >= 100


def test_no_torch_types_in_abi():
    src = open(os.path.join(ROOT, "include", "b200ep.h")).read()
    assert "torch" not in src.lower().replace("torch.distributed", "").replace("torch.bool", "")
    assert "at::" not in src and "c10::" not in src


def test_size_hint_and_errors_without_gpu():
    from b200ep import _lib

    lib = _lib.load()
    n = lib.b200ep_ll_size_hint(256, 7168, 8, 256)
    assert 1 << 30 < n < 8 << 30
    assert lib.b200ep_ll_size_hint(256, 7168, 8, 257) == 0  # experts must divide over ranks
    import ctypes as C

    ctx = C.c_void_p()
    rc = lib.b200ep_create(C.byref(ctx), 9, 8, 0, 1 << 30)  # rank out of range: rejected before any CUDA call
    assert rc == 1 and b"bad rank/world" in lib.b200ep_last_error()
    with pytest.raises(_lib.B200EPError):
        _lib.check(rc, "create")


def test_product_path_does_not_import_the_oracle():
    """The oracle is test infrastructure: nothing under llm-d_b200/ may touch it."""
    pkg = os.path.join(ROOT, "llm-d_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f"{f} imports oracle"
