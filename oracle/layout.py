"""Top-k routing layout, CPU restatement (test infrastructure).

Restates what ``deep_ep.Buffer.get_dispatch_layout`` returns to vLLM
(``prepare_finalize/deepep_ht.py:121-133``): per-rank and per-expert token
counts and the token x rank membership matrix, for experts sharded linearly,
``owner(e) = e // (E / W)`` (vLLM ``fused_moe/expert_map_manager.py:75-79``).
The counting half is the same arithmetic as the head of ``cpu_fused_moe_torch``
(``cpu_fused_moe.py:444-446``: ``cnts.scatter_(1, topk_ids, 1); cnts.sum(0)``).

``send_off`` is this library's addition: the exclusive prefix count of the
membership column, i.e. the position a token takes among the tokens its rank
sends to that destination.  It is what makes the received order canonical
(source rank ascending, source token ascending) without atomics.
Negative ids are "no expert" and are skipped, as DeepEP does.
"""

from __future__ import annotations

import numpy as np


def expert_owner(ids: np.ndarray, num_experts: int, world: int) -> np.ndarray:
    assert num_experts % world == 0
    return np.where(ids >= 0, ids // (num_experts // world), -1)


def dispatch_layout(topk_ids: np.ndarray, num_experts: int, world: int):
    """Returns (tokens_per_rank i32[W], tokens_per_expert i32[E],
    is_token_in_rank u8[T,W], send_off i32[T,W])."""
    ids = np.asarray(topk_ids).astype(np.int64)
    T, K = ids.shape
    valid = ids >= 0
    tokens_per_expert = np.bincount(ids[valid], minlength=num_experts).astype(np.int32)
    owner = expert_owner(ids, num_experts, world)
    in_rank = np.zeros((T, world), np.uint8)
    for k in range(K):
        rows = np.nonzero(valid[:, k])[0]
        in_rank[rows, owner[rows, k]] = 1
    tokens_per_rank = in_rank.sum(axis=0).astype(np.int32)
    incl = np.cumsum(in_rank.astype(np.int64), axis=0)
    send_off = np.where(in_rank == 1, incl - 1, -1).astype(np.int32)
    return tokens_per_rank, tokens_per_expert, in_rank, send_off
