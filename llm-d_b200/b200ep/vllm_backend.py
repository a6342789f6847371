"""vLLM-facing classes that use b200ep's fused quantise+dispatch (SURVEY.md Route B).

Route A (``llm-d_b200/deep_ep``) already lets vLLM's own DeepEP manager and prepare/finalize
classes run on b200ep unchanged.  What Route A cannot do is skip the separate FP8 quantisation
pass vLLM runs before the high-throughput dispatch
(vllm/model_executor/layers/fused_moe/prepare_finalize/deepep_ht.py:282-289): the class below
overrides just that step.  Everything else -- receiver, finalize, handles per DBO micro-batch,
the manager (vllm/distributed/device_communicators/all2all.py:143-325) -- is inherited.

``install()`` swaps the class in where vLLM constructs it
(vllm/model_executor/layers/fused_moe/all2all_utils.py:155); call it from a
``vllm.general_plugins`` entry point (vllm/plugins/__init__.py:14,69-82) so that every worker
process does it.  The llm-d manifests keep ``--all2all-backend deepep_high_throughput``.
"""
from __future__ import annotations

from typing import Callable

import torch

_classes = None


def get_classes():
    """Build the subclasses lazily (needs vLLM and the deep_ep shim importable)."""
    global _classes
    if _classes is not None:
        return _classes

    import deep_ep  # the b200ep shim (llm-d_b200/deep_ep)
    import vllm.model_executor.layers.fused_moe.modular_kernel as mk
    from vllm.model_executor.layers.fused_moe.config import FusedMoEQuantConfig
    from vllm.model_executor.layers.fused_moe.prepare_finalize.deepep_ht import DeepEPHTPrepareAndFinalize
    from vllm.v1.worker.ubatching import (dbo_current_ubatch_id, dbo_enabled, dbo_get_previous_event,
                                          dbo_switch_to_compute_sync, dbo_yield_and_switch_from_compute_to_comm)

    def _ue8m0_wanted() -> bool:
        try:
            from vllm.utils.deep_gemm import is_deep_gemm_e8m0_used

            return bool(is_deep_gemm_e8m0_used())
        except Exception:
            return False

    class B200EPHTPrepareAndFinalize(DeepEPHTPrepareAndFinalize):
        """DeepEPHTPrepareAndFinalize with the FP8 block quantisation folded into the dispatch."""

        @staticmethod
        def can_fuse(buffer, a1: torch.Tensor, quant_config: FusedMoEQuantConfig, defer_input_quant: bool) -> bool:
            bs = quant_config.block_shape
            return (isinstance(buffer, deep_ep.Buffer) and hasattr(buffer, "_impl") and not defer_input_quant
                    and quant_config.is_block_quantized and quant_config.quant_dtype == torch.float8_e4m3fn
                    and bs is not None and bs[1] == 128 and not quant_config.per_act_token_quant
                    and a1.dtype == torch.bfloat16 and a1.size(-1) % 128 == 0)

        def prepare_async(self, a1: torch.Tensor, topk_weights: torch.Tensor, topk_ids: torch.Tensor, num_experts: int,
                          expert_map: torch.Tensor | None, apply_router_weight_on_input: bool,
                          quant_config: FusedMoEQuantConfig, defer_input_quant: bool = False) -> mk.ReceiverType:
            if not self.can_fuse(self.buffer, a1, quant_config, defer_input_quant):
                return super().prepare_async(a1, topk_weights, topk_ids, num_experts, expert_map,
                                             apply_router_weight_on_input, quant_config, defer_input_quant)
            if apply_router_weight_on_input:
                assert topk_ids.size(1) == 1, "apply_router_weight_on_input is only implemented for topk=1"
                a1 = a1 * topk_weights.to(a1.dtype)
            return self._do_dispatch_fused(a1, topk_ids, topk_weights, num_experts, quant_config)

        def _do_dispatch_fused(self, tokens, rank_topk_ids, rank_topk_weights, num_experts, quant_config) -> Callable:
            # same sequence as DeepEPHTPrepareAndFinalize._do_dispatch (deepep_ht.py:97-181)
            previous_event = dbo_get_previous_event(self.buffer.capture)
            dbo_yield_and_switch_from_compute_to_comm()
            (num_tokens_per_rank, num_tokens_per_rdma_rank, dispatch_expert_num_tokens, is_token_in_rank,
             event) = self.buffer.get_dispatch_layout(topk_idx=rank_topk_ids, num_experts=num_experts,
                                                      previous_event=previous_event, async_finish=False,
                                                      allocate_on_comm_stream=False)
            (token_data, expert_topk_ids, expert_topk_weights, expert_num_tokens_per_expert_list, handle,
             event) = self.buffer.dispatch(
                x=tokens, handle=None, num_tokens_per_rank=num_tokens_per_rank,
                num_tokens_per_rdma_rank=num_tokens_per_rdma_rank, is_token_in_rank=is_token_in_rank,
                num_tokens_per_expert=dispatch_expert_num_tokens, topk_idx=rank_topk_ids,
                topk_weights=rank_topk_weights, expert_alignment=1, config=self._get_dispatch_config(),
                previous_event=previous_event, async_finish=self.async_prepare and not dbo_enabled(),
                allocate_on_comm_stream=False, b200ep_quantize=True, b200ep_ue8m0=_ue8m0_wanted())
            self.handles[dbo_current_ubatch_id()] = handle
            dbo_switch_to_compute_sync()
            return lambda: self._receiver(event, True, token_data, expert_topk_ids, num_experts,
                                          expert_num_tokens_per_expert_list, expert_topk_weights, None, quant_config,
                                          defer_input_quant=False)

    class B200EPHTNoSyncPrepareAndFinalize(B200EPHTPrepareAndFinalize):
        """Host-sync-free high-throughput prepare: the dispatch never waits on the host, the
        received tensors have a fixed worst-case number of rows (``max_tokens_per_rank * num_dispatchers``) and the
        per-expert counts stay on the device, so ``ExpertTokensMetadata.expert_num_tokens_cpu`` is None -- the form
        vLLM's experts already accept from the low-latency path (deepep_ll.py:328-342) and the reason it has to turn
        CUDA graphs off for ``deepep_high_throughput`` today (config/compilation.py:1189-1205, deepep_ht.py:222-227).
        Rows beyond the received count hold stale data with expert id -1 / weight 0 semantics: their top-k ids are
        set to the out-of-rank sentinel, so experts that honour expert_map skip them."""

        def __init__(self, *args, max_tokens_per_rank: int, **kwargs):
            super().__init__(*args, **kwargs)
            self.max_tokens_per_rank_ = int(max_tokens_per_rank)

        def _do_dispatch_fused(self, tokens, rank_topk_ids, rank_topk_weights, num_experts, quant_config) -> Callable:
            assert tokens.shape[0] <= self.max_tokens_per_rank_, "more tokens than the worst case this object was built for"
            previous_event = dbo_get_previous_event(self.buffer.capture)
            dbo_yield_and_switch_from_compute_to_comm()
            worst = self.max_tokens_per_rank_ * self.num_dispatchers_
            (token_data, expert_topk_ids, expert_topk_weights, _, handle, event) = self.buffer.dispatch(
                x=tokens, handle=None, num_tokens_per_rank=None, num_tokens_per_rdma_rank=None, is_token_in_rank=None,
                num_tokens_per_expert=torch.empty(num_experts, dtype=torch.int32, device="meta"), topk_idx=rank_topk_ids,
                topk_weights=rank_topk_weights, expert_alignment=1, num_worst_tokens=worst,
                config=self._get_dispatch_config(), previous_event=previous_event,
                async_finish=self.async_prepare and not dbo_enabled(), allocate_on_comm_stream=False,
                b200ep_quantize=True, b200ep_ue8m0=_ue8m0_wanted())
            self.handles[dbo_current_ubatch_id()] = handle
            dbo_switch_to_compute_sync()

            def receiver():
                if event.event is not None:
                    event.current_stream_wait()
                expert_x, expert_x_scale = token_data
                rows = torch.arange(worst, device=expert_x.device)
                live = (rows < handle.num_recv_dev).unsqueeze(1)
                sentinel = num_experts - 1 if self.rank_expert_offset == 0 else 0
                ids = torch.where(live & (expert_topk_ids != -1), expert_topk_ids + self.rank_expert_offset, sentinel)
                weights = torch.where(live, expert_topk_weights, 0.0)
                meta = mk.ExpertTokensMetadata(expert_num_tokens=handle.recv_per_expert_dev, expert_num_tokens_cpu=None)
                return expert_x, expert_x_scale, meta, ids, weights

            return receiver

    _classes = {"B200EPHTPrepareAndFinalize": B200EPHTPrepareAndFinalize,
                "B200EPHTNoSyncPrepareAndFinalize": B200EPHTNoSyncPrepareAndFinalize}
    return _classes


# ---------------------------------------------------------------------------------------------
# Expert-grouped layout without the Triton passes.  Drop-in replacements, same
# signatures, for the two functions DeepGemmExperts.apply calls around its grouped GEMMs
# (vllm/model_executor/layers/fused_moe/experts/deep_gemm_moe.py:279-291, 322-329; defined in
# fused_moe/deep_gemm_utils.py:347-440).  Rows that did not come from a b200ep dispatch fall through to
# vLLM's own functions.


def deepgemm_moe_permute(aq, aq_scale, topk_ids, local_num_experts, expert_map, expert_tokens_meta, aq_out=None):
    """``deepgemm_moe_permute`` on the rows a b200ep dispatch received: one plan kernel + one scatter kernel
    (b200ep_ht_permute) instead of ep_scatter's two Triton kernels, the fill of expert_ids and the count pass;
    rows inside an expert come out in received order (deterministic)."""
    from vllm.model_executor.layers.fused_moe import deep_gemm_utils as dgu

    from .buffer import _RECV_REGISTRY

    ent = _RECV_REGISTRY.get(aq.data_ptr())
    if ent is None or tuple(aq.shape) != (ent[1].num_recv, ent[1].H) or aq.dtype != torch.float8_e4m3fn:
        return dgu.deepgemm_moe_permute(aq, aq_scale, topk_ids, local_num_experts, expert_map, expert_tokens_meta, aq_out)
    buf, h = ent
    block_m = dgu.get_mk_alignment_for_contiguous_layout()[0]
    per_expert = None
    if expert_tokens_meta is not None and expert_tokens_meta.expert_num_tokens_cpu is not None:
        per_expert = expert_tokens_meta.expert_num_tokens_cpu.tolist()
    a_perm, s_perm, expert_ids, inv_perm = buf.ht_permute(h, alignment=block_m, per_expert=per_expert, out_x=aq_out)
    return a_perm, s_perm, expert_ids, inv_perm


def deepgemm_unpermute_and_reduce(a, topk_ids, topk_weights, inv_perm, expert_map, output):
    """``deepgemm_unpermute_and_reduce``, deferred: when ``inv_perm`` came from the function above, nothing runs here.
    The grouped GEMM output stays where it is and the combine that follows (deep_ep.Buffer.combine on the handle of
    the same dispatch) gathers, weights and reduces inside its push (b200ep_ht_combine_permuted); ``output`` is not
    written, and the combine ignores it."""
    from vllm.model_executor.layers.fused_moe import deep_gemm_utils as dgu

    from .buffer import _RECV_REGISTRY

    for buf, h in list(_RECV_REGISTRY.values()):
        if tuple(inv_perm.shape) == (h.num_recv, h.K) and inv_perm.dtype == torch.int32 and getattr(h, "_inv_ptr", 0) == inv_perm.data_ptr():
            w = topk_weights.to(torch.float32).contiguous()
            h.deferred = (a, inv_perm, w)
            return None
    return dgu.deepgemm_unpermute_and_reduce(a, topk_ids, topk_weights, inv_perm, expert_map, output)


def install() -> None:
    """vLLM general-plugin hook: make maybe_make_prepare_finalize build the fused class, and let the DeepGEMM experts
    regroup received rows / return results through b200ep instead of ep_scatter / ep_gather."""
    import vllm.model_executor.layers.fused_moe.all2all_utils as a2a

    a2a.DeepEPHTPrepareAndFinalize = get_classes()["B200EPHTPrepareAndFinalize"]
    try:
        import vllm.model_executor.layers.fused_moe.experts.deep_gemm_moe as dgm

        dgm.deepgemm_moe_permute = deepgemm_moe_permute
        dgm.deepgemm_unpermute_and_reduce = deepgemm_unpermute_and_reduce
    except Exception:  # the experts module needs more than this backend does; the P&F swap above still stands
        pass
