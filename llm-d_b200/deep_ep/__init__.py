"""``deep_ep``-compatible surface over b200ep (SURVEY.md Route A).

vLLM selects its expert-parallel all2all by the literal backend names
``deepep_high_throughput`` / ``deepep_low_latency`` and finds the kernels with
``importlib.util.find_spec("deep_ep")`` (vllm/utils/import_utils.py:394-405).  Putting
this directory on ``PYTHONPATH`` therefore makes the llm-d wide-EP manifests
(guides/wide-ep-lws/modelserver/gpu/vllm/base/{decode.yaml:119,prefill.yaml:91}) run on
b200ep unchanged.  Only the part of DeepEP's Python API that vLLM 0.22/0.23 touches is
provided -- exactly the list in SURVEY.md Appendix A:

  Buffer(group, num_nvl_bytes, num_rdma_bytes, low_latency_mode, num_qps_per_rank, ...)
      vllm/distributed/device_communicators/all2all.py:222-229,297-306
  Buffer.set_num_sms / get_low_latency_rdma_size_hint / get_dispatch_config / get_combine_config
  buf.capture / get_dispatch_layout / dispatch / combine
      vllm/model_executor/layers/fused_moe/prepare_finalize/deepep_ht.py:114-162,362-376
  buf.low_latency_dispatch / low_latency_combine
      vllm/model_executor/layers/fused_moe/prepare_finalize/deepep_ll.py:292-314,399-408
  buf.destroy, EventOverlap.event / current_stream_wait

Differences from DeepEP that a caller can observe:
  * recv tensors are zero-copy views of the symmetric region; they stay valid until the
    second next dispatch on the same Buffer (vLLM consumes them well before that);
  * the received order is deterministic (source rank, then source token);
  * the region actually allocated is max(requested bytes, what the shapes need, see
    ``B200EP_HT_BYTES``): B200 has 180 GB, the 1 GiB DeepEP default is only a floor.
"""
from __future__ import annotations

import os
from typing import Callable

import torch

import b200ep
from b200ep import _lib as _L

__version__ = "1.2.1+b200ep"
__all__ = ["Buffer", "Config", "EventOverlap"]


class EventOverlap:
    """DeepEP's event wrapper: ``.event`` may be None (nothing to wait for)."""

    def __init__(self, event: torch.cuda.Event | None = None, extra_tensors=None):
        self.event = event
        self.extra_tensors = extra_tensors  # keeps tensors alive across the async window

    def current_stream_wait(self) -> None:
        if self.event is not None:
            torch.cuda.current_stream().wait_event(self.event)

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.current_stream_wait()


class Config:
    """Opaque tuning object (DeepEP: num_sms + chunk sizes).  b200ep sizes its grids itself;
    only num_sms is honoured, through Buffer.set_num_sms."""

    def __init__(self, num_sms: int = 20, *chunk_sizes):
        self.num_sms = num_sms
        self.chunk_sizes = chunk_sizes


class Buffer:
    num_sms: int = 0  # 0 = let b200ep use every SM (DeepEP default is 20 of a Hopper's 132)

    def __init__(self, group, num_nvl_bytes: int = 0, num_rdma_bytes: int = 0, low_latency_mode: bool = False,
                 num_qps_per_rank: int = 1, allow_nvlink_for_low_latency_mode: bool = True, allow_mnnvl: bool = False,
                 explicitly_destroy: bool = False, **_ignored):
        self.group = group
        self.low_latency_mode = low_latency_mode
        self.explicitly_destroy = explicitly_destroy
        want = max(int(num_nvl_bytes), int(num_rdma_bytes))
        if not low_latency_mode:
            want = max(want, int(os.environ.get("B200EP_HT_BYTES", 4 << 30)))
        self._impl = b200ep.Buffer(group, sym_bytes=max(want, 64 << 20))
        self.rank, self.group_size = self._impl.rank, self._impl.world
        self._comm = torch.cuda.Stream(device=self._impl.device)
        if Buffer.num_sms:
            self._impl.set_num_sms(Buffer.num_sms)

    # ------------------------------------------------------------ static configuration
    @staticmethod
    def set_num_sms(new_num_sms: int) -> None:
        Buffer.num_sms = int(new_num_sms)

    @staticmethod
    def get_low_latency_rdma_size_hint(num_max_dispatch_tokens_per_rank: int, hidden: int, num_ranks: int,
                                       num_experts: int) -> int:
        return b200ep.Buffer.ll_size_hint(num_max_dispatch_tokens_per_rank, hidden, num_ranks, num_experts)

    @staticmethod
    def get_dispatch_config(num_ranks: int) -> Config:
        return Config(Buffer.num_sms or 20)

    @staticmethod
    def get_combine_config(num_ranks: int) -> Config:
        return Config(Buffer.num_sms or 20)

    def destroy(self) -> None:
        torch.cuda.synchronize()
        self._impl.destroy()

    def __del__(self):
        if not getattr(self, "explicitly_destroy", True):
            try:
                self.destroy()
            except Exception:
                pass

    # ------------------------------------------------------------ stream plumbing
    def capture(self) -> EventOverlap:
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        return EventOverlap(ev)

    def _enter_comm(self, previous_event: EventOverlap | None):
        compute = torch.cuda.current_stream()
        if previous_event is not None and previous_event.event is not None:
            self._comm.wait_event(previous_event.event)
        else:
            self._comm.wait_stream(compute)
        return compute

    def _leave_comm(self, compute, async_finish: bool, tensors) -> EventOverlap:
        if async_finish:
            ev = torch.cuda.Event()
            ev.record(self._comm)
            for t in tensors:
                if isinstance(t, torch.Tensor) and t.is_cuda:
                    t.record_stream(self._comm)
            return EventOverlap(ev, tensors)
        compute.wait_stream(self._comm)
        return EventOverlap(None)

    # ------------------------------------------------------------ high throughput
    def get_dispatch_layout(self, topk_idx: torch.Tensor, num_experts: int, previous_event: EventOverlap | None = None,
                            async_finish: bool = False, allocate_on_comm_stream: bool = False):
        compute = self._enter_comm(previous_event)
        with torch.cuda.stream(self._comm):
            lay = self._impl.get_dispatch_layout(topk_idx.contiguous(), num_experts)
        ev = self._leave_comm(compute, async_finish, (topk_idx,))
        # dispatch() recomputes the layout inside its single C call (a few microseconds of GPU time): nothing
        # has to be smuggled along on these tensors, and the caller may clone or slice them freely
        return lay.tokens_per_rank, None, lay.tokens_per_expert, lay.is_token_in_rank, ev

    def dispatch(self, x, handle=None, num_tokens_per_rank=None, num_tokens_per_rdma_rank=None, is_token_in_rank=None,
                 num_tokens_per_expert=None, topk_idx=None, topk_weights=None, expert_alignment: int = 1,
                 num_worst_tokens: int = 0, config: Config | None = None, previous_event: EventOverlap | None = None,
                 async_finish: bool = False, allocate_on_comm_stream: bool = False, b200ep_quantize: bool = False,
                 b200ep_ue8m0: bool = False):
        """DeepEP signature plus two b200ep extensions: ``b200ep_quantize=True`` with a bf16 ``x``
        quantises to e4m3 / 128-group fp32 scales INSIDE the dispatch kernel (what vLLM otherwise
        does with per_token_group_quant_fp8 before the call, deepep_ht.py:282-289) and returns the
        (recv_x, recv_scales) tuple; ``b200ep_ue8m0`` rounds those scales up to powers of two."""
        if handle is not None:
            raise NotImplementedError("b200ep: cached-handle re-dispatch is not used by vLLM and not provided")
        assert topk_idx is not None, "b200ep dispatch needs topk_idx"
        if isinstance(x, tuple):
            xq, xs = x
            xs = xs.contiguous()
        else:
            xq, xs = x, None
        assert num_tokens_per_expert is not None, "b200ep dispatch needs num_tokens_per_expert (its length is the expert count)"
        num_experts = int(num_tokens_per_expert.numel())
        compute = self._enter_comm(previous_event)
        with torch.cuda.stream(self._comm):
            # one C call: layout kernel (its last block publishes the counts) -> dispatch kernel (waits for the
            # peers' counts itself) -> host wait for this rank's row counts, which runs under the dispatch kernel
            rx, rs, ridx, rw, per_expert, h = self._impl.dispatch(
                xq.contiguous(), topk_idx.contiguous(), None if topk_weights is None else topk_weights.contiguous(),
                num_experts, x_scales=xs, quantize=b200ep_quantize, ue8m0=b200ep_ue8m0,
                num_worst_tokens=num_worst_tokens)
        ev = self._leave_comm(compute, async_finish, (xq, xs, topk_idx, topk_weights))
        if expert_alignment > 1:
            per_expert = [(n + expert_alignment - 1) // expert_alignment * expert_alignment for n in per_expert]
        recv_x = (rx, rs) if rs is not None else rx
        return recv_x, ridx, rw, per_expert, h, ev

    def combine(self, x: torch.Tensor, handle, topk_weights=None, bias=None, config: Config | None = None,
                previous_event: EventOverlap | None = None, async_finish: bool = False,
                allocate_on_comm_stream: bool = False):
        if topk_weights is not None or bias is not None:
            raise NotImplementedError("b200ep: combine(topk_weights=..., bias=...) is not used by vLLM (deepep_ht.py:367-376)")
        compute = self._enter_comm(previous_event)
        deferred, handle.deferred = getattr(handle, "deferred", None), None
        with torch.cuda.stream(self._comm):
            if deferred is not None:
                # the experts left their grouped GEMM output in place (b200ep.vllm_backend.deepgemm_unpermute_and_reduce):
                # gather + weighted reduction happen inside the push, `x` was never written
                y_grouped, inv_perm, weights = deferred
                out = self._impl.combine_permuted(y_grouped, inv_perm, handle, topk_weights=weights)
            else:
                out = self._impl.combine(x.contiguous(), handle)
        ev = self._leave_comm(compute, async_finish, (x, out))
        return out, None, ev

    # ------------------------------------------------------------ low latency
    def low_latency_dispatch(self, x: torch.Tensor, topk_idx: torch.Tensor, num_max_dispatch_tokens_per_rank: int,
                             num_experts: int, cumulative_local_expert_recv_stats=None, use_fp8: bool = True,
                             round_scale: bool = False, use_ue8m0: bool = False, async_finish: bool = False,
                             return_recv_hook: bool = False, **unsupported):
        if unsupported.get("use_nvfp4"):
            raise NotImplementedError("b200ep: NVFP4 low-latency dispatch is not implemented (SURVEY.md 8(f) f3); "
                                      "run vLLM with VLLM_DEEPEPLL_NVFP4_DISPATCH=0 (bf16 dispatch, its default)")
        # use_ue8m0: packed exponent bytes (one per 128 values, four per int32, slot index fastest -- what DeepGEMM
        # reads on sm_100, vLLM fp8_utils.py:643-714), written by the dispatch kernel itself; like the payload
        # they are complete only after hook() when return_recv_hook is set.  round_scale alone: fp32 powers of two.
        mode = 2 if (use_fp8 and use_ue8m0) else (1 if round_scale else 0)
        px, ps, cnt, h, hook = self._impl.low_latency_dispatch(
            x.contiguous(), topk_idx.contiguous(), num_max_dispatch_tokens_per_rank, num_experts, use_fp8=use_fp8,
            ue8m0=mode, return_recv_hook=return_recv_hook)
        recv_x = (px, ps) if use_fp8 else px
        ev = EventOverlap(self.capture().event if async_finish else None)
        return recv_x, cnt, h, ev, (hook if return_recv_hook else None)

    def low_latency_combine(self, x: torch.Tensor, topk_idx: torch.Tensor, topk_weights: torch.Tensor, handle,
                            use_logfmt: bool = False, zero_copy: bool = False, async_finish: bool = False,
                            return_recv_hook: bool = False, out: torch.Tensor | None = None, combine_wait_recv_cost_stats=None):
        o, hook = self._impl.low_latency_combine(x.contiguous(), topk_idx.contiguous(),
                                                 topk_weights.to(torch.float32).contiguous(), handle, out,
                                                 return_recv_hook=return_recv_hook)
        ev = EventOverlap(self.capture().event if async_finish else None)
        return o, ev, (hook if return_recv_hook else None)

    def clean_low_latency_buffer(self, *a, **k) -> None:  # epochs make explicit cleaning unnecessary
        return None
