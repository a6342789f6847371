"""The reference's CPU dispatch/combine, ported to a standalone torch function
(test infrastructure; also the thing timed by ``bench.py``'s ``cpu_baseline`` and
``--impl reference`` legs -- kind "port").

It follows vLLM ``model_executor/layers/fused_moe/cpu_fused_moe.py`` function
``cpu_fused_moe_torch``:

  :444-446  cnts.scatter_(1, topk_ids, 1); tokens_per_expert = cnts.sum(0)
  :447      idxs = topk_ids.view(-1).argsort()          (unstable on CPU)
  :449      sorted_tokens = input[idxs // K]            <- DISPATCH (index_select)
  :455-465  per-expert MLP over contiguous slices       <- replaced by ``expert_fn``
  :467-470  new_x = empty_like(outs); new_x[idxs] = outs  <- COMBINE scatter-back
  :474-480  new_x.view(T,K,H).type(f32).mul_(w[..., None]).sum(1).type(bf16)

``expert_fn(e, rows)`` stands in for the expert MLP; the default is the identity,
which is what the benchmarks use (the GEMMs are out of scope).
``oracle/gen_golden.py`` checks this port against the real ``cpu_fused_moe_torch``
imported from the installed vLLM and stores the result as a golden fixture.
"""

from __future__ import annotations

import torch


def cpu_dispatch(input: torch.Tensor, topk_ids: torch.Tensor, num_experts: int, stable: bool = False):
    cnts = topk_ids.new_zeros((topk_ids.shape[0], num_experts))
    cnts.scatter_(1, topk_ids.to(torch.int64), 1)
    tokens_per_expert = cnts.sum(dim=0)
    idxs = topk_ids.view(-1).argsort(stable=stable) if stable else topk_ids.view(-1).argsort()
    sorted_tokens = input[idxs // topk_ids.shape[1]]
    return sorted_tokens, tokens_per_expert, idxs


def cpu_combine(outs: torch.Tensor, idxs: torch.Tensor, topk_weights: torch.Tensor, topk_shape):
    new_x = torch.empty_like(outs)
    new_x[idxs] = outs
    return (
        new_x.view(*topk_shape, -1)
        .type(topk_weights.dtype)
        .mul_(topk_weights.unsqueeze(dim=-1))
        .sum(dim=1)
        .type(new_x.dtype)
    )


def cpu_dispatch_combine(
    input: torch.Tensor,
    topk_weights: torch.Tensor,
    topk_ids: torch.Tensor,
    num_experts: int,
    expert_fn=None,
    stable: bool = False,
) -> torch.Tensor:
    sorted_tokens, tokens_per_expert, idxs = cpu_dispatch(input, topk_ids, num_experts, stable)
    if expert_fn is None:
        outs = sorted_tokens
    else:
        outputs = []
        start = 0
        for e, n in enumerate(tokens_per_expert.tolist()):
            if n == 0:
                continue
            outputs.append(expert_fn(e, sorted_tokens[start : start + n]))
            start += n
        outs = torch.cat(outputs, dim=0) if outputs else sorted_tokens.new_empty(0)
    return cpu_combine(outs, idxs, topk_weights, topk_ids.shape)
