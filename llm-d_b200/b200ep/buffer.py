"""Host side of b200ep: one ``Buffer`` per rank over the C ABI (``include/b200ep.h``).

The class mirrors the operations vLLM performs on ``deep_ep.Buffer``
(vllm/distributed/device_communicators/all2all.py:222-245,297-315 and
vllm/model_executor/layers/fused_moe/prepare_finalize/deepep_{ht,ll}.py) with
native names; ``llm-d_b200/deep_ep`` wraps it in DeepEP's exact signatures.
torch is used for device memory, streams and the bootstrap all-gather only;
every kernel is reached through ctypes with raw pointers.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Any, Sequence

import torch

from . import _lib as L

DEFAULT_SYM_BYTES = 4 << 30


class _DevMem:
    """A raw device range exposed through __cuda_array_interface__ (zero copy)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {
            "shape": (nbytes,),
            "typestr": "|u1",
            "data": (ptr, False),
            "version": 3,
            "strides": None,
        }


def _ptr(t: torch.Tensor | None) -> int | None:
    return None if t is None else t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


@dataclass
class Layout:
    tokens_per_rank: torch.Tensor    # int32 [W]
    tokens_per_expert: torch.Tensor  # int32 [E]
    is_token_in_rank: torch.Tensor   # bool  [T, W]
    send_off: torch.Tensor           # int32 [T, W]
    num_experts: int


@dataclass
class HtHandle:
    slot: int
    T: int
    H: int
    K: int
    E: int
    x_kind: int
    layout: Layout
    num_recv: int = -1
    rank_prefix: list[int] = field(default_factory=list)
    keep: Any = None  # tensors that must outlive the asynchronous kernels
    num_recv_dev: Any = None         # host-sync-free dispatch: rows received, device int32 scalar
    recv_per_expert_dev: Any = None  # host-sync-free dispatch: per-local-expert counts, device int32 [E/W]
    deferred: Any = None             # (y_grouped, inv_perm, weights) left by a deferred unpermute: the combine gathers itself


@dataclass
class LlHandle:
    slot: int
    T: int
    H: int
    K: int
    E: int
    max_tok: int
    use_fp8: bool
    packed_ue8m0: bool = False  # scales come back as packed exponent bytes (int32, slot index fastest)
    src_info: torch.Tensor | None = None
    layout_range: torch.Tensor | None = None
    keep: Any = None


class _HtSlotState:
    __slots__ = ("key", "lay", "lay_ptrs", "el", "n", "n_ref", "pe", "rp", "full", "vcache", "num_recv_dev",
                 "recv_per_expert_dev")

    def __init__(self, key, lay: Layout, el: int, world: int):
        self.key, self.lay, self.el = key, lay, el
        self.lay_ptrs = (lay.tokens_per_rank.data_ptr(), lay.tokens_per_expert.data_ptr(),
                         lay.is_token_in_rank.data_ptr(), lay.send_off.data_ptr())
        self.n = C.c_int()
        self.n_ref = C.byref(self.n)
        self.pe = (C.c_int32 * el)()
        self.rp = (C.c_int32 * (world + 1))()
        self.full = None
        self.vcache = {}
        self.num_recv_dev = self.recv_per_expert_dev = None

    def views(self, M: int):
        """(recv_x, recv_scales, recv_topk_idx, recv_topk_weights, recv_src_idx) narrowed to M rows."""
        v = self.vcache.get(M)
        if v is None:
            fx, fs, fi, fw, fr = self.full[0]
            assert 0 <= M <= self.full[2]
            if len(self.vcache) > 64:
                self.vcache.clear()
            v = self.vcache[M] = (fx[:M], (None if fs is None else fs[:M]), fi[:M], fw[:M], fr[:M])
        return v


# recv_x.data_ptr() -> (Buffer, HtHandle) of the latest dispatch on that slot: lets code that only sees the received
# tensor (vLLM's expert classes) find the dispatch it came from (b200ep.vllm_backend.deepgemm_moe_permute)
_RECV_REGISTRY: dict = {}


def exchange_handles(group, mine: bytes) -> bytes:
    """All-gather one 64-byte CUDA-IPC handle per rank over a host process group (gloo or nccl);
    returns the world*64-byte blob in rank order.  Same pattern as vLLM's custom all-reduce
    bootstrap (vllm/distributed/device_communicators/custom_all_reduce.py:296-315)."""
    import torch.distributed as dist

    assert len(mine) == 64
    world = dist.get_world_size(group)
    gathered: list = [None] * world
    dist.all_gather_object(gathered, mine, group=group)
    assert all(isinstance(g, bytes) and len(g) == 64 for g in gathered)
    return b"".join(gathered)


class Buffer:
    """One rank's endpoint.  ``group`` is any torch.distributed group (gloo or nccl);
    it is used once, to all-gather the 64-byte CUDA-IPC handles, and for barriers."""

    def __init__(self, group=None, *, sym_bytes: int = DEFAULT_SYM_BYTES, device: int | None = None,
                 rank: int | None = None, world: int | None = None, _connect: bool = True):
        self.lib = L.load()
        if not torch.cuda.is_available():
            raise L.B200EPError("b200ep needs a CUDA device; there is no CPU fallback")
        import torch.distributed as dist

        self.group = group
        if group is not None:
            self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        else:
            self.rank, self.world = (rank or 0), (world or 1)
        self.device = torch.cuda.current_device() if device is None else device
        torch.cuda.set_device(self.device)
        torch.cuda.current_stream().synchronize()  # make sure a CUDA context exists
        ctx = C.c_void_p()
        L.check(self.lib.b200ep_create(C.byref(ctx), self.rank, self.world, self.device, sym_bytes), "b200ep_create")
        self.ctx = ctx
        base, nbytes = C.c_void_p(), C.c_size_t()
        L.check(self.lib.b200ep_sym_base(self.ctx, C.byref(base), C.byref(nbytes)))
        self.sym_ptr, self.sym_bytes = base.value, nbytes.value
        self._sym = torch.as_tensor(_DevMem(self.sym_ptr, self.sym_bytes), device=f"cuda:{self.device}")
        self._slot = 0
        self._destroyed = False
        if _connect and self.world > 1:
            self._connect_ipc()

    # ------------------------------------------------------------ bootstrap
    def _connect_ipc(self):
        import torch.distributed as dist

        h = (C.c_char * 64)()
        L.check(self.lib.b200ep_export_handle(self.ctx, h), "export_handle")
        blob = exchange_handles(self.group, bytes(h.raw))
        L.check(self.lib.b200ep_import_handles(self.ctx, blob, self.world), "import_handles")
        dist.barrier(group=self.group)

    @staticmethod
    def connect_local(buffers: Sequence["Buffer"]):
        """Wire several Buffers living in ONE process (tests / single-GPU emulation)."""
        n = len(buffers)
        arr = (C.c_void_p * n)(*[b.sym_ptr for b in buffers])
        for b in buffers:
            assert b.world == n
            L.check(b.lib.b200ep_set_peer_ptrs(b.ctx, arr, n), "set_peer_ptrs")

    def destroy(self):
        if not self._destroyed:
            self._destroyed = True
            for k in [k for k, v in _RECV_REGISTRY.items() if v[0] is self]:
                del _RECV_REGISTRY[k]
            self._sym = None
            L.check(self.lib.b200ep_destroy(self.ctx), "destroy")

    def set_num_sms(self, n: int):
        L.check(self.lib.b200ep_set_num_sms(self.ctx, int(n)))

    def _view(self, ptr: int, nbytes: int) -> torch.Tensor:
        off = ptr - self.sym_ptr
        assert 0 <= off and off + nbytes <= self.sym_bytes
        return self._sym[off : off + nbytes]

    def next_slot(self) -> int:
        s = self._slot
        self._slot = (s + 1) % L.NUM_SLOTS
        return s

    # ------------------------------------------------------------ layout
    def get_dispatch_layout(self, topk_idx: torch.Tensor, num_experts: int) -> Layout:
        assert topk_idx.is_cuda and topk_idx.dim() == 2 and topk_idx.is_contiguous()
        assert topk_idx.dtype in (torch.int64, torch.int32)
        T, K = topk_idx.shape
        dev = topk_idx.device
        tpr = torch.empty(self.world, dtype=torch.int32, device=dev)
        tpe = torch.empty(num_experts, dtype=torch.int32, device=dev)
        inr = torch.empty((T, self.world), dtype=torch.bool, device=dev)
        off = torch.empty((T, self.world), dtype=torch.int32, device=dev)
        L.check(
            self.lib.b200ep_layout(self.ctx, _ptr(topk_idx), int(topk_idx.dtype == torch.int64), T, K, num_experts,
                                   _ptr(tpr), _ptr(tpe), _ptr(inr), _ptr(off), _stream()),
            "b200ep_layout",
        )
        return Layout(tpr, tpe, inr, off, num_experts)

    @staticmethod
    def quant_fp8(x: torch.Tensor, ue8m0: bool = False):
        """Stand-alone 128-group e4m3 quantiser (same device code as the fused dispatch)."""
        assert x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 2 and x.is_contiguous()
        T, H = x.shape
        q = torch.empty((T, H), dtype=torch.float8_e4m3fn, device=x.device)
        s = torch.empty((T, H // 128), dtype=torch.float32, device=x.device)
        L.check(L.load().b200ep_quant_fp8(_ptr(x), _ptr(q), _ptr(s), T, H, int(ue8m0), _stream()), "quant_fp8")
        return q, s

    # ------------------------------------------------------------ high throughput
    def ht_capacity(self, H: int, K: int, x_kind: int) -> int:
        return self.lib.b200ep_ht_capacity(self.ctx, H, K, x_kind)

    def ht_notify(self, slot: int, layout: Layout, phases: int = L.PHASE_ALL):
        L.check(
            self.lib.b200ep_ht_notify(self.ctx, slot, _ptr(layout.tokens_per_rank), _ptr(layout.tokens_per_expert),
                                      int(layout.is_token_in_rank.shape[0]), layout.num_experts, phases, _stream()),
            "ht_notify",
        )

    def ht_notify_wait(self, slot: int, num_experts: int, timeout_ms: int = 120000):
        el = num_experts // self.world
        n = C.c_int()
        pe = (C.c_int32 * el)()
        rp = (C.c_int32 * (self.world + 1))()
        L.check(self.lib.b200ep_ht_notify_wait(self.ctx, slot, C.byref(n), pe, rp, timeout_ms), "ht_notify_wait")
        return n.value, list(pe), list(rp)

    def ht_dispatch_launch(self, h: HtHandle, x, x_scales, topk_idx, topk_weights, ue8m0: bool, phases: int):
        L.check(
            self.lib.b200ep_ht_dispatch(
                self.ctx, h.slot, _ptr(x), _ptr(x_scales), h.x_kind, int(ue8m0), _ptr(topk_idx),
                int(topk_idx.dtype == torch.int64), _ptr(topk_weights), _ptr(h.layout.is_token_in_rank),
                _ptr(h.layout.send_off), h.T, h.H, h.K, h.E, phases, _stream()),
            "ht_dispatch",
        )

    def ht_views(self, h: HtHandle):
        """Zero-copy tensors over this rank's receive region (valid until the slot is reused,
        i.e. until the second next dispatch on this Buffer)."""
        M, H, K = h.num_recv, h.H, h.K
        st = self.__dict__.get("_ht_state", {}).get(h.slot)
        if st is not None and st.full is not None and st.key[1:5] == (H, K, h.E, h.x_kind):
            return st.views(M)
        # full-capacity views are built once per (slot, shape): the region layout only depends on them;
        # a call then costs five narrow() ops instead of ~20 tensor ops (the EP=1 step is host bound)
        cache = self.__dict__.setdefault("_view_cache", {})
        key = (h.slot, H, K, h.x_kind)
        full = cache.get(h.slot)
        if full is None or full[0] != key:
            v = L.HtView()
            L.check(self.lib.b200ep_ht_views(self.ctx, h.slot, C.byref(v)), "ht_views")
            cap = v.capacity_rows
            if h.x_kind == L.X_BF16:
                fx = self._view(v.recv_x, cap * 2 * H).view(torch.bfloat16).view(cap, H)
                fs = None
            else:
                fx = self._view(v.recv_x, cap * H).view(torch.float8_e4m3fn).view(cap, H)
                fs = self._view(v.recv_x_scales, cap * (H // 128) * 4).view(torch.float32).view(cap, H // 128)
            fi = self._view(v.recv_topk_idx, cap * K * 8).view(torch.int64).view(cap, K)
            fw = self._view(v.recv_topk_weights, cap * K * 4).view(torch.float32).view(cap, K)
            fr = self._view(v.recv_src_idx, cap * 4).view(torch.int32)
            full = (key, (fx, fs, fi, fw, fr), cap)
            cache[h.slot] = full
        fx, fs, fi, fw, fr = full[1]
        assert 0 <= M <= full[2]
        return fx[:M], (None if fs is None else fs[:M]), fi[:M], fw[:M], fr[:M]

    def dispatch(self, x, topk_idx, topk_weights, num_experts: int, *, x_scales=None, quantize: bool = False,
                 ue8m0: bool = False, layout: Layout | None = None, num_worst_tokens: int = 0):
        """Whole high-throughput dispatch (collective).  x: bf16 [T,H] (moved as is, or
        quantised in the kernel when ``quantize``), or e4m3 [T,H] with ``x_scales``.
        Returns (recv_x, recv_x_scales, recv_topk_idx, recv_topk_weights,
        num_recv_tokens_per_expert_list, handle).

        ``num_worst_tokens > 0`` selects the host-sync-free form (DeepEP's parameter of the same name;
        SURVEY.md 8(f) f2): nothing waits on the host, the views have ``num_worst_tokens`` rows of which
        only the first ``handle.num_recv_dev`` (a device scalar) are meaningful, the per-expert list is
        empty and ``handle.recv_per_expert_dev`` holds the counts on the device."""
        if layout is not None:
            h, args = self._dispatch_prepare(x, topk_idx, topk_weights, num_experts, x_scales, quantize, ue8m0, layout)
            self.ht_notify(h.slot, h.layout)
            per_expert = self._dispatch_after_notify(h)
            self.ht_dispatch_launch(h, *args, L.PHASE_ALL)
            views = self.ht_views(h)
            _RECV_REGISTRY[views[0].data_ptr()] = (self, h)
            return (*views[:4], per_expert, h)
        # no layout given: layout + count exchange + payload in ONE C call (b200ep_ht_dispatch_all)
        T, H = x.shape
        K = topk_idx.shape[1]
        if x.dtype == torch.bfloat16:
            kind = L.X_BF16_QUANT if quantize else L.X_BF16
        elif x.dtype == torch.float8_e4m3fn:
            kind = L.X_FP8
            assert x_scales is not None and x_scales.dtype == torch.float32 and x_scales.is_contiguous()
        else:
            raise L.B200EPError(f"unsupported dispatch dtype {x.dtype}")
        assert x.is_cuda and x.is_contiguous() and topk_idx.is_contiguous() and topk_idx.shape[0] == T
        assert topk_weights is None or (topk_weights.dtype == torch.float32 and topk_weights.is_contiguous())
        slot = self.next_slot()
        st = self._ht_slot_state(slot, T, H, K, num_experts, kind, x.device)
        nosync = num_worst_tokens > 0
        rc = self.lib.b200ep_ht_dispatch_all(
            self.ctx, slot, x.data_ptr(), _ptr(x_scales), kind, int(ue8m0), topk_idx.data_ptr(),
            int(topk_idx.dtype == torch.int64), _ptr(topk_weights), st.lay_ptrs[0], st.lay_ptrs[1], st.lay_ptrs[2],
            st.lay_ptrs[3], T, H, K, num_experts, st.n_ref, st.pe, st.rp, -1 if nosync else 120000, _stream())
        if rc:
            L.check(rc, "ht_dispatch_all")
        keep = (x, x_scales, topk_idx, topk_weights)
        if st.full is None:
            self._ht_slot_views(st, slot, H, K, kind)
        if nosync:
            if num_worst_tokens > st.full[2]:
                raise L.B200EPError(f"num_worst_tokens={num_worst_tokens} exceeds the slot capacity of {st.full[2]} rows")
            h = HtHandle(slot, T, H, K, num_experts, kind, st.lay, int(num_worst_tokens), [], keep)
            h.num_recv_dev, h.recv_per_expert_dev = st.num_recv_dev, st.recv_per_expert_dev
            views = st.views(h.num_recv)
            _RECV_REGISTRY[views[0].data_ptr()] = (self, h)
            return (*views[:4], [], h)
        h = HtHandle(slot, T, H, K, num_experts, kind, st.lay, st.n.value, st.rp[:], keep)
        views = st.views(h.num_recv)
        _RECV_REGISTRY[views[0].data_ptr()] = (self, h)
        return (*views[:4], st.pe[:], h)

    def _ht_slot_state(self, slot, T, H, K, E, kind, dev):
        """Everything a dispatch on `slot` needs that does not change from call to call while the shape stays
        the same: the layout tensors (kept until the matching combine), the ctypes outputs of the C call, the
        full-capacity views of the receive region.  The EP=1 step is bound by this host code, not by the GPU."""
        cache = self.__dict__.setdefault("_ht_state", {})
        key = (T, H, K, E, kind, str(dev))
        st = cache.get(slot)
        if st is None or st.key != key:
            st = _HtSlotState(key, self._layout_tensors(slot, T, E, dev), E // self.world, self.world)
            cache[slot] = st
        return st

    def _ht_slot_views(self, st, slot, H, K, kind):
        v = L.HtView()
        L.check(self.lib.b200ep_ht_views(self.ctx, slot, C.byref(v)), "ht_views")
        cap = v.capacity_rows
        if kind == L.X_BF16:
            fx = self._view(v.recv_x, cap * 2 * H).view(torch.bfloat16).view(cap, H)
            fs = None
        else:
            fx = self._view(v.recv_x, cap * H).view(torch.float8_e4m3fn).view(cap, H)
            fs = self._view(v.recv_x_scales, cap * (H // 128) * 4).view(torch.float32).view(cap, H // 128)
        fi = self._view(v.recv_topk_idx, cap * K * 8).view(torch.int64).view(cap, K)
        fw = self._view(v.recv_topk_weights, cap * K * 4).view(torch.float32).view(cap, K)
        fr = self._view(v.recv_src_idx, cap * 4).view(torch.int32)
        st.full = ((fx, fs, fi, fw, fr), None, cap)
        devs = self._sym.device
        prefix = torch.as_tensor(_DevMem(v.rank_prefix, (self.world + 1) * 4), device=devs).view(torch.int32)
        st.num_recv_dev = prefix[self.world]
        st.recv_per_expert_dev = torch.as_tensor(_DevMem(v.recv_per_expert, st.el * 4), device=devs).view(torch.int32)

    def _layout_tensors(self, slot: int, T: int, E: int, dev) -> Layout:
        """Layout outputs are needed until the matching combine; keep one set per slot and reuse it
        while the shape stays the same (saves four allocator calls per dispatch)."""
        cache = self.__dict__.setdefault("_lay_cache", {})
        key = (slot, T, E, str(dev))
        lay = cache.get(slot)
        if lay is None or lay[0] != key:
            lay = (key, Layout(torch.empty(self.world, dtype=torch.int32, device=dev),
                               torch.empty(E, dtype=torch.int32, device=dev),
                               torch.empty((T, self.world), dtype=torch.bool, device=dev),
                               torch.empty((T, self.world), dtype=torch.int32, device=dev), E))
            cache[slot] = lay
        return lay[1]

    def _dispatch_prepare(self, x, topk_idx, topk_weights, num_experts, x_scales, quantize, ue8m0, layout):
        assert x.is_cuda and x.dim() == 2 and x.is_contiguous()
        assert topk_idx.is_contiguous() and topk_idx.shape[0] == x.shape[0]
        if x.dtype == torch.bfloat16:
            kind = L.X_BF16_QUANT if quantize else L.X_BF16
            assert x_scales is None
        elif x.dtype == torch.float8_e4m3fn:
            kind = L.X_FP8
            assert x_scales is not None and x_scales.dtype == torch.float32 and x_scales.is_contiguous()
            assert tuple(x_scales.shape) == (x.shape[0], x.shape[1] // 128)
        else:
            raise L.B200EPError(f"unsupported dispatch dtype {x.dtype}")
        if topk_weights is not None:
            assert topk_weights.dtype == torch.float32 and topk_weights.is_contiguous()
        if layout is None:
            layout = self.get_dispatch_layout(topk_idx, num_experts)
        T, H = x.shape
        h = HtHandle(self.next_slot(), T, H, topk_idx.shape[1], num_experts, kind, layout)
        h.keep = (x, x_scales, topk_idx, topk_weights)
        return h, (x, x_scales, topk_idx, topk_weights, ue8m0)

    def _dispatch_after_notify(self, h: HtHandle):
        h.num_recv, per_expert, h.rank_prefix = self.ht_notify_wait(h.slot, h.E)
        return per_expert

    def combine(self, y: torch.Tensor, handle: HtHandle, out: torch.Tensor | None = None, phases: int = L.PHASE_ALL):
        """High-throughput combine (collective).  y: bf16 [num_recv, H] in received order."""
        h = handle
        assert y.dtype == torch.bfloat16 and y.is_contiguous() and tuple(y.shape) == (h.num_recv, h.H), (
            f"combine expects bf16 [{h.num_recv}, {h.H}], got {y.dtype} {tuple(y.shape)}")
        if out is None:
            out = torch.empty((h.T, h.H), dtype=torch.bfloat16, device=y.device)
        assert out.dtype == torch.bfloat16 and out.is_contiguous() and tuple(out.shape) == (h.T, h.H)
        L.check(
            self.lib.b200ep_ht_combine(self.ctx, h.slot, _ptr(y), _ptr(out), _ptr(h.layout.is_token_in_rank),
                                       _ptr(h.layout.send_off), h.T, h.H, phases, _stream()),
            "ht_combine",
        )
        return out

    # ------------------------------------------------------------ expert-grouped layout (DeepGEMM contiguous)
    @staticmethod
    def aligned_rows(num_rows: int, K: int, local_experts: int, alignment: int = 128, per_expert=None) -> int:
        """Rows the grouped layout needs: the exact sum of the per-expert counts rounded up to ``alignment`` when
        the counts are known on the host, else the worst case (vLLM ``compute_aligned_M``,
        fused_moe/deep_gemm_utils.py:26-44)."""
        if per_expert is not None:
            return sum((int(n) + alignment - 1) // alignment * alignment for n in per_expert)
        m = num_rows * K + local_experts * (alignment - 1)
        return (m + alignment - 1) // alignment * alignment

    def ht_permute(self, handle: HtHandle, *, alignment: int = 128, per_expert=None, out_x=None, out_scales=None):
        """Regroup the rows ``handle``'s dispatch received by local expert (vLLM ``deepgemm_moe_permute``,
        fused_moe/deep_gemm_utils.py:347-420).  Returns (a_perm [M_sum, H], a_scale_perm [M_sum, H/128] or None,
        expert_ids [M_sum] int32, inv_perm [num_recv, K] int32).  ``per_expert`` (the list the dispatch returned)
        sizes the outputs exactly; without it (host-sync-free dispatch) they are sized for the worst case."""
        h = handle
        el = h.E // self.world
        dev = self._sym.device
        m_cap = self.aligned_rows(h.num_recv, h.K, el, alignment, per_expert if per_expert else None)
        bf16 = h.x_kind == L.X_BF16
        if out_x is None:
            out_x = torch.empty((m_cap, h.H), dtype=torch.bfloat16 if bf16 else torch.float8_e4m3fn, device=dev)
        assert out_x.is_contiguous() and out_x.shape[0] >= m_cap and out_x.shape[1] == h.H
        if not bf16 and out_scales is None:
            out_scales = torch.empty((m_cap, h.H // 128), dtype=torch.float32, device=dev)
        expert_ids = torch.empty(m_cap, dtype=torch.int32, device=dev)
        inv_perm = torch.empty((h.num_recv, h.K), dtype=torch.int32, device=dev)
        L.check(
            self.lib.b200ep_ht_permute(self.ctx, h.slot, alignment, h.num_recv, _ptr(out_x), _ptr(out_scales),
                                       _ptr(expert_ids), _ptr(inv_perm), None, m_cap, _stream()),
            "ht_permute",
        )
        h._inv_ptr = inv_perm.data_ptr()  # lets a deferred unpermute recognise its own inv_perm (vllm_backend)
        return out_x[:m_cap], (None if bf16 else out_scales[:m_cap]), expert_ids, inv_perm

    def combine_permuted(self, y_grouped: torch.Tensor, inv_perm: torch.Tensor, handle: HtHandle, out=None, *,
                         topk_weights: torch.Tensor | None = None, phases: int = L.PHASE_ALL):
        """``deepgemm_unpermute_and_reduce`` + ``Buffer.combine`` in one operation: the weighted fp32 reduction over a
        received token's local experts is done inside the push (see b200ep_ht_combine_permuted).
        y_grouped: bf16 [M_sum, H], rows as laid out by ``ht_permute``."""
        h = handle
        assert y_grouped.dtype == torch.bfloat16 and y_grouped.is_contiguous() and y_grouped.shape[1] == h.H
        assert inv_perm.dtype == torch.int32 and inv_perm.is_contiguous() and tuple(inv_perm.shape) == (h.num_recv, h.K)
        if topk_weights is not None:
            assert topk_weights.dtype == torch.float32 and topk_weights.is_contiguous()
            assert tuple(topk_weights.shape) == (h.num_recv, h.K)
        if out is None:
            out = torch.empty((h.T, h.H), dtype=torch.bfloat16, device=y_grouped.device)
        assert out.dtype == torch.bfloat16 and out.is_contiguous() and tuple(out.shape) == (h.T, h.H)
        L.check(
            self.lib.b200ep_ht_combine_permuted(self.ctx, h.slot, _ptr(y_grouped), _ptr(inv_perm), _ptr(topk_weights),
                                                _ptr(out), _ptr(h.layout.is_token_in_rank), _ptr(h.layout.send_off),
                                                h.T, h.H, phases, _stream()),
            "ht_combine_permuted",
        )
        return out

    # ------------------------------------------------------------ low latency
    @staticmethod
    def ll_size_hint(max_tokens_per_rank: int, hidden: int, world: int, num_experts: int) -> int:
        return L.load().b200ep_ll_size_hint(max_tokens_per_rank, hidden, world, num_experts)

    def ll_dispatch_launch(self, h: LlHandle, x, topk_idx, ue8m0: bool, phases: int):
        L.check(
            self.lib.b200ep_ll_dispatch(self.ctx, h.slot, _ptr(x), _ptr(topk_idx), int(topk_idx.dtype == torch.int64),
                                        h.T, h.H, h.K, h.E, h.max_tok, int(h.use_fp8), int(ue8m0), phases, _stream()),
            "ll_dispatch",
        )

    def ll_views(self, h: LlHandle):
        v = L.LlView()
        L.check(self.lib.b200ep_ll_views(self.ctx, h.slot, C.byref(v)), "ll_views")
        el, cap, H = h.E // self.world, self.world * h.max_tok, h.H
        if h.use_fp8 and h.packed_ue8m0:
            # int32 [el, g4, cap4] in memory; handed out as [el, cap, g4] with the slot index fastest -- the
            # "MN-major, TMA-aligned" layout DeepGEMM wants for packed UE8M0 scales (vLLM fp8_utils.py:643-714)
            g4, cap4 = (H // 128 + 3) // 4, (cap + 3) // 4 * 4
            px = self._view(v.packed_recv_x, el * cap * H).view(torch.float8_e4m3fn).view(el, cap, H)
            ps = self._view(v.packed_recv_scales, el * g4 * cap4 * 4).view(torch.int32).view(el, g4, cap4)
            ps = ps.permute(0, 2, 1)[:, :cap, :]
        elif h.use_fp8:
            px = self._view(v.packed_recv_x, el * cap * H).view(torch.float8_e4m3fn).view(el, cap, H)
            ps = self._view(v.packed_recv_scales, el * cap * (H // 128) * 4).view(torch.float32).view(el, cap, H // 128)
        else:
            px = self._view(v.packed_recv_x, el * cap * H * 2).view(torch.bfloat16).view(el, cap, H)
            ps = None
        cnt = torch.as_tensor(_DevMem(v.recv_count, el * 4), device=self._sym.device).view(torch.int32)
        src = self._view(v.src_info, el * cap * 4).view(torch.int32).view(el, cap)
        rng = torch.as_tensor(_DevMem(v.layout_range, el * self.world * 8), device=self._sym.device).view(
            torch.int32).view(el, self.world, 2)
        return px, ps, cnt, src, rng

    def low_latency_dispatch(self, x, topk_idx, max_tokens_per_rank: int, num_experts: int, *, use_fp8: bool = True,
                             ue8m0: bool | int = False, return_recv_hook: bool = False):
        """Returns (packed_recv_x, packed_recv_scales, recv_count, handle, hook).
        ``ue8m0``: False/0 = fp32 scales, True/1 = fp32 scales rounded up to powers of two,
        2 = the same scales as packed exponent bytes (int32 view [E_local, slots, H/512], slot index fastest),
        written by the kernel itself -- like every other output they are complete once the hook has run."""
        assert x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 2 and x.is_contiguous()
        assert topk_idx.is_contiguous() and topk_idx.shape[0] == x.shape[0]
        T, H = x.shape
        ue8m0 = int(ue8m0)
        h = LlHandle(self.next_slot(), T, H, topk_idx.shape[1], num_experts, max_tokens_per_rank, use_fp8,
                     packed_ue8m0=(use_fp8 and ue8m0 == 2))
        h.keep = (x, topk_idx)
        first = L.LL_PHASE_COUNT | L.LL_PHASE_SEND
        self.ll_dispatch_launch(h, x, topk_idx, ue8m0, first if return_recv_hook else L.LL_PHASE_ALL)
        px, ps, cnt, h.src_info, h.layout_range = self.ll_views(h)
        hook = (lambda: self.ll_dispatch_launch(h, x, topk_idx, ue8m0, L.LL_PHASE_RECV)) if return_recv_hook else None
        return px, ps, cnt, h, hook

    def low_latency_combine(self, y, topk_idx, topk_weights, handle: LlHandle, out=None, *,
                            return_recv_hook: bool = False, phases: int | None = None):
        h = handle
        el, cap = h.E // self.world, self.world * h.max_tok
        assert y.dtype == torch.bfloat16 and y.is_contiguous() and tuple(y.shape) == (el, cap, h.H)
        assert topk_weights.dtype == torch.float32 and topk_weights.is_contiguous()
        if out is None:
            out = torch.empty((h.T, h.H), dtype=torch.bfloat16, device=y.device)
        assert out.dtype == torch.bfloat16 and out.is_contiguous() and tuple(out.shape) == (h.T, h.H)

        def launch(ph):
            L.check(
                self.lib.b200ep_ll_combine(self.ctx, h.slot, _ptr(y), _ptr(topk_idx),
                                           int(topk_idx.dtype == torch.int64), _ptr(topk_weights), _ptr(out), h.T, h.H,
                                           h.K, h.E, h.max_tok, ph, _stream()),
                "ll_combine",
            )

        if phases is not None:
            launch(phases)
            return out, None
        if return_recv_hook:
            launch(L.PHASE_SEND)
            return out, (lambda: launch(L.PHASE_RECV))
        launch(L.PHASE_ALL)
        return out, None


class LocalWorld:
    """W ranks emulated in ONE process on ONE GPU: each collective runs phase by phase
    over all ranks (no kernel ever waits for one launched later), as the profiling guide
    prescribes for boxes with fewer GPUs than ranks.  Used by the single-GPU parity tests
    and smoke(); the production path is one process per GPU over CUDA IPC."""

    def __init__(self, world: int, sym_bytes: int = 1 << 30, device: int | None = None):
        self.world = world
        self.bufs = [Buffer(None, sym_bytes=sym_bytes, device=device, rank=r, world=world, _connect=False)
                     for r in range(world)]
        if world > 1:
            Buffer.connect_local(self.bufs)

    def destroy(self):
        torch.cuda.synchronize()
        for b in self.bufs:
            b.destroy()

    def dispatch(self, xs, topk_idx, topk_weights, num_experts, *, x_scales=None, quantize=False, ue8m0=False):
        W = self.world
        sc = x_scales if x_scales is not None else [None] * W
        prep = [self.bufs[r]._dispatch_prepare(xs[r], topk_idx[r], topk_weights[r], num_experts, sc[r], quantize,
                                               ue8m0, None) for r in range(W)]
        for r in range(W):
            self.bufs[r].ht_notify(prep[r][0].slot, prep[r][0].layout, L.PHASE_SEND)
        for r in range(W):
            self.bufs[r].ht_notify(prep[r][0].slot, prep[r][0].layout, L.PHASE_RECV)
        per_expert = [self.bufs[r]._dispatch_after_notify(prep[r][0]) for r in range(W)]
        for r in range(W):
            self.bufs[r].ht_dispatch_launch(prep[r][0], *prep[r][1], L.PHASE_SEND)
        for r in range(W):
            self.bufs[r].ht_dispatch_launch(prep[r][0], *prep[r][1], L.PHASE_RECV)
        return [(*self.bufs[r].ht_views(prep[r][0]), per_expert[r], prep[r][0]) for r in range(W)]

    def combine(self, ys, handles, outs=None):
        W = self.world
        res = []
        for r in range(W):
            o = None if outs is None else outs[r]
            h = handles[r]
            if o is None:
                o = torch.empty((h.T, h.H), dtype=torch.bfloat16, device=ys[r].device)
            res.append(o)
            self.bufs[r].combine(ys[r], h, o, L.PHASE_SEND)
        for r in range(W):
            self.bufs[r].combine(ys[r], handles[r], res[r], L.PHASE_RECV)
        return res

    def combine_permuted(self, ys_grouped, inv_perms, handles, topk_weights=None):
        W = self.world
        outs = [torch.empty((handles[r].T, handles[r].H), dtype=torch.bfloat16, device=ys_grouped[r].device)
                for r in range(W)]
        for ph in (L.PHASE_SEND, L.PHASE_RECV):
            for r in range(W):
                self.bufs[r].combine_permuted(ys_grouped[r], inv_perms[r], handles[r], outs[r],
                                              topk_weights=None if topk_weights is None else topk_weights[r], phases=ph)
        return outs

    def low_latency_dispatch(self, xs, topk_idx, max_tok, num_experts, *, use_fp8=True, ue8m0=False):
        W = self.world
        hs = []
        for r in range(W):
            T, H = xs[r].shape
            hs.append(LlHandle(self.bufs[r].next_slot(), T, H, topk_idx[r].shape[1], num_experts, max_tok, use_fp8,
                               packed_ue8m0=(use_fp8 and int(ue8m0) == 2)))
        for ph in (L.LL_PHASE_COUNT, L.LL_PHASE_SEND, L.LL_PHASE_RECV):
            for r in range(W):
                self.bufs[r].ll_dispatch_launch(hs[r], xs[r], topk_idx[r], ue8m0, ph)
        out = []
        for r in range(W):
            px, ps, cnt, hs[r].src_info, hs[r].layout_range = self.bufs[r].ll_views(hs[r])
            out.append((px, ps, cnt, hs[r]))
        return out

    def low_latency_combine(self, ys, topk_idx, topk_weights, handles):
        W = self.world
        outs = [torch.empty((handles[r].T, handles[r].H), dtype=torch.bfloat16, device=ys[r].device) for r in range(W)]
        for ph in (L.PHASE_SEND, L.PHASE_RECV):
            for r in range(W):
                self.bufs[r].low_latency_combine(ys[r], topk_idx[r], topk_weights[r], handles[r], outs[r], phases=ph)
        return outs
