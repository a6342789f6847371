"""ctypes binding of libb200ep.so (the C ABI declared in include/b200ep.h).

There is no CPU fallback: if the CUDA library is missing or a call fails this
module raises.  Build it with ``python -c "import __graft_entry__ as g; g.build()"``
or ``make -C llm-d_b200/csrc``.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200ep.so")

MAX_RANKS = 8
MAX_LOCAL_EXPERTS = 1024
MAX_TOPK = 16
NUM_SLOTS = 2

X_BF16, X_BF16_QUANT, X_FP8 = 0, 1, 2
PHASE_SEND, PHASE_RECV, PHASE_ALL = 1, 2, 3
LL_PHASE_COUNT, LL_PHASE_SEND, LL_PHASE_RECV, LL_PHASE_ALL = 1, 2, 4, 7


class B200EPError(RuntimeError):
    pass


class HtView(C.Structure):
    _fields_ = [
        ("recv_x", C.c_void_p),
        ("recv_x_scales", C.c_void_p),
        ("recv_topk_idx", C.c_void_p),
        ("recv_topk_weights", C.c_void_p),
        ("recv_src_idx", C.c_void_p),
        ("rank_prefix", C.c_void_p),
        ("recv_per_expert", C.c_void_p),
        ("capacity_rows", C.c_int32),
    ]


class LlView(C.Structure):
    _fields_ = [
        ("packed_recv_x", C.c_void_p),
        ("packed_recv_scales", C.c_void_p),
        ("recv_count", C.c_void_p),
        ("src_info", C.c_void_p),
        ("layout_range", C.c_void_p),
    ]


_p, _i, _sz = C.c_void_p, C.c_int, C.c_size_t
# name -> (restype, argtypes); must list every symbol of include/b200ep.h (tests/test_abi.py checks)
SIGNATURES = {
    "b200ep_last_error": (C.c_char_p, []),
    "b200ep_version": (_i, []),
    "b200ep_create": (_i, [C.POINTER(_p), _i, _i, _i, _sz]),
    "b200ep_destroy": (_i, [_p]),
    "b200ep_export_handle": (_i, [_p, _p]),
    "b200ep_import_handles": (_i, [_p, _p, _i]),
    "b200ep_set_peer_ptrs": (_i, [_p, C.POINTER(_p), _i]),
    "b200ep_sym_base": (_i, [_p, C.POINTER(_p), C.POINTER(_sz)]),
    "b200ep_peer_base": (_i, [_p, _i, C.POINTER(_p)]),
    "b200ep_debug_trace": (_i, [_p, _p]),
    "b200ep_set_num_sms": (_i, [_p, _i]),
    "b200ep_layout": (_i, [_p, _p, _i, _i, _i, _i, _p, _p, _p, _p, _p]),
    "b200ep_layout_noalloc": (_i, [_i, _p, _i, _i, _i, _i, _p, _p, _p, _p, _p]),
    "b200ep_quant_fp8": (_i, [_p, _p, _p, _i, _i, _i, _p]),
    "b200ep_ht_capacity": (_i, [_p, _i, _i, _i]),
    "b200ep_ht_notify": (_i, [_p, _i, _p, _p, _i, _i, _i, _p]),
    "b200ep_ht_notify_wait": (_i, [_p, _i, C.POINTER(_i), _p, _p, _i]),
    "b200ep_ht_dispatch": (_i, [_p, _i, _p, _p, _i, _i, _p, _i, _p, _p, _p, _i, _i, _i, _i, _i, _p]),
    "b200ep_ht_dispatch_all": (_i, [_p, _i, _p, _p, _i, _i, _p, _i, _p, _p, _p, _p, _p, _i, _i, _i, _i, C.POINTER(_i), _p,
                                    _p, _i, _p]),
    "b200ep_ht_views": (_i, [_p, _i, C.POINTER(HtView)]),
    "b200ep_ht_combine": (_i, [_p, _i, _p, _p, _p, _p, _i, _i, _i, _p]),
    "b200ep_ht_permute": (_i, [_p, _i, _i, _i, _p, _p, _p, _p, _p, _i, _p]),
    "b200ep_ht_combine_permuted": (_i, [_p, _i, _p, _p, _p, _p, _p, _p, _i, _i, _i, _p]),
    "b200ep_ll_size_hint": (_sz, [_i, _i, _i, _i]),
    "b200ep_ll_dispatch": (_i, [_p, _i, _p, _p, _i, _i, _i, _i, _i, _i, _i, _i, _i, _p]),
    "b200ep_ll_views": (_i, [_p, _i, C.POINTER(LlView)]),
    "b200ep_ll_combine": (_i, [_p, _i, _p, _p, _i, _p, _p, _i, _i, _i, _i, _i, _i, _p]),
}

_lib = None


def load():
    """Load the shared library (no CUDA call is made).  Raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B200EPError(
            f"{LIB_PATH} is missing: build the CUDA library first (make -C llm-d_b200/csrc). "
            "b200ep has no CPU fallback."
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so does not export it
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = load().b200ep_last_error().decode(errors="replace")
        raise B200EPError(f"{what or 'b200ep'} failed (code {rc}): {msg}")
