"""Combine (gather back + top-k reduction), CPU restatement (test infrastructure).

Arithmetic being restated
-------------------------
``cpu_fused_moe_torch`` scatters expert outputs back to (token, k) order and
reduces ``new_x.view(T, K, H).type(float32).mul_(w.unsqueeze(-1)).sum(1)
.type(bf16)`` (vLLM ``cpu_fused_moe.py:467-481``): an fp32 weighted sum over the
top-k entries, rounded once to bf16.  Across ranks the same sum is split by
owner: ``AgRsAll2AllManager.combine`` reduce-scatters per-rank partials
(``all2all.py:124-137``).

* High-throughput: vLLM applies the weights and the within-rank sum on the
  expert rank first (``deepep_ht.py:350-361``, ``topk_weight_and_reduce.py:80-121``),
  then ``deep_ep.Buffer.combine(x=bf16 [M,H], topk_weights=None)``
  (``deepep_ht.py:364-376``) adds the <= W partial rows of a token.  North-star
  arithmetic: fp32 accumulate, one bf16 rounding.  Order: destination rank
  ascending (deterministic, and the order the CUDA kernel uses, so the two agree
  bit for bit).
* Low-latency: ``low_latency_combine`` does the weighting itself,
  ``out[t] = sum_k w[t,k] * y[e_k, slot(t,k)]`` (``deepep_ll.py:391-410``), fp32
  multiply then fp32 add (no fused multiply-add), k ascending, one bf16 rounding.
"""

from __future__ import annotations

import numpy as np

from .bits import bf16_to_f32, f32_to_bf16


def ht_combine(ys, disp, tokens_per_src):
    """ys[d]: [M_d, H] bf16 bits produced on expert rank d in received order.
    disp: the list ``oracle.dispatch.ht_dispatch`` returned.  tokens_per_src[s]: T_s.
    Returns out[s]: [T_s, H] bf16 bits."""
    W = len(ys)
    H = np.asarray(ys[0]).shape[1]
    acc = [np.zeros((tokens_per_src[s], H), np.float32) for s in range(W)]
    for d in range(W):  # ascending destination rank == accumulation order
        y = bf16_to_f32(np.asarray(ys[d]))
        pre = disp[d]["rank_prefix"]
        src = disp[d]["recv_src_idx"]
        for s in range(W):
            a, b = int(pre[s]), int(pre[s + 1])
            if b > a:
                t = src[a:b]
                acc[s][t] = (acc[s][t] + y[a:b]).astype(np.float32)
    return [f32_to_bf16(a) for a in acc]


def ll_combine(ys, topk_ids, topk_weights, disp, num_experts: int):
    """ys[d]: [E_local, W*max_tok, H] bf16 bits (expert outputs in the packed
    layout).  Returns out[s]: [T_s, H] bf16 bits."""
    W = len(ys)
    E_local = num_experts // W
    H = np.asarray(ys[0]).shape[2]
    outs = []
    for s in range(W):
        ids = np.asarray(topk_ids[s]).astype(np.int64)
        w = np.asarray(topk_weights[s], np.float32)
        T, K = ids.shape
        # slot of every (t, k): invert src_info of the owning rank
        slot = np.full((T, K), -1, np.int64)
        for d in range(W):
            rng = disp[d]["layout_range"]
            info = disp[d]["src_info"]
            for el in range(E_local):
                a, n = int(rng[el, s, 0]), int(rng[el, s, 1])
                flat = info[el, a : a + n]
                slot[flat // K, flat % K] = a + np.arange(n)
        acc = np.zeros((T, H), np.float32)
        for k in range(K):
            for t in range(T):
                e = ids[t, k]
                if e < 0:
                    continue
                d, el = divmod(int(e), E_local)
                row = bf16_to_f32(np.asarray(ys[d])[el, slot[t, k]])
                prod = (row * w[t, k]).astype(np.float32)
                acc[t] = (acc[t] + prod).astype(np.float32)
        outs.append(f32_to_bf16(acc))
    return outs


def ht_combine_source(num_tokens: int, hidden: int, parts):
    """High-throughput combine as ONE source rank sees it (same arithmetic as ``ht_combine``).
    parts: one ``(token_idx, y_rows)`` pair per destination rank, ASCENDING rank order: the
    source tokens that rank received from this source (ascending) and the bf16 rows it sends
    back for them.  Returns out [num_tokens, hidden] bf16 bits."""
    acc = np.zeros((num_tokens, hidden), np.float32)
    for tok, rows in parts:
        tok = np.asarray(tok, np.int64)
        if tok.size:
            acc[tok] = (acc[tok] + bf16_to_f32(np.asarray(rows))).astype(np.float32)
    return f32_to_bf16(acc)


def ll_combine_source(topk_ids, topk_weights, rows, num_experts: int):
    """Low-latency combine as ONE source rank sees it (same arithmetic as ``ll_combine``).
    rows: [T, K, H] bf16 bits, rows[t, k] = the expert output row for entry (t, k) (ignored
    where the id is negative or >= num_experts).  fp32 multiply, then fp32 add, k ascending."""
    ids = np.asarray(topk_ids).astype(np.int64)
    w = np.asarray(topk_weights, np.float32)
    T, K = ids.shape
    acc = np.zeros((T, np.asarray(rows).shape[2]), np.float32)
    for k in range(K):
        m = (ids[:, k] >= 0) & (ids[:, k] < num_experts)
        prod = (bf16_to_f32(np.asarray(rows)[:, k]) * w[:, k][:, None]).astype(np.float32)
        acc[m] = (acc[m] + prod[m]).astype(np.float32)
    return f32_to_bf16(acc)
