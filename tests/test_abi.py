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


def test_ctypes_signatures_match_header_parameter_counts():
    """Every prototype in include/b200ep.h has as many parameters as its ctypes argtypes entry."""
    from b200ep import _lib

    src = open(os.path.join(ROOT, "include", "b200ep.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    protos = re.findall(r"\b(b200ep_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", src, flags=re.S)
    assert len(protos) == len(_lib.SIGNATURES)
    for name, params in protos:
        params = params.strip()
        n = 0 if params in ("", "void") else params.count(",") + 1
        assert n == len(_lib.SIGNATURES[name][1]), f"{name}: header has {n} parameters, ctypes {len(_lib.SIGNATURES[name][1])}"


def test_vllm_deepep_modules_import_against_the_shim():
    """Route A: with llm-d_b200 on the path, `deep_ep` resolves to the shim and the two vLLM modules
    that import it (and nothing else in vLLM does) load -- no GPU needed for that."""
    pytest.importorskip("vllm")
    import importlib

    import deep_ep

    assert deep_ep.__file__.startswith(os.path.join(ROOT, "llm-d_b200"))
    for attr in ("Buffer", "Config", "EventOverlap"):
        assert hasattr(deep_ep, attr)
    for meth in ("get_dispatch_layout", "dispatch", "combine", "low_latency_dispatch", "low_latency_combine", "capture",
                 "destroy", "set_num_sms", "get_low_latency_rdma_size_hint", "get_dispatch_config", "get_combine_config"):
        assert hasattr(deep_ep.Buffer, meth), meth
    from vllm.utils.import_utils import has_deep_ep

    assert has_deep_ep()
    ht = importlib.import_module("vllm.model_executor.layers.fused_moe.prepare_finalize.deepep_ht")
    ll = importlib.import_module("vllm.model_executor.layers.fused_moe.prepare_finalize.deepep_ll")
    assert ht.DeepEPHTPrepareAndFinalize and ll.DeepEPLLPrepareAndFinalize
    from b200ep import vllm_backend

    cls = vllm_backend.get_classes()["B200EPHTPrepareAndFinalize"]
    assert issubclass(cls, ht.DeepEPHTPrepareAndFinalize)
