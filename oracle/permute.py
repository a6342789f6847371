"""Expert-grouped permute of received rows and its inverse with the weighted reduction, CPU restatement
(test infrastructure).

What is restated
----------------
vLLM's DeepGEMM expert path regroups the rows a high-throughput dispatch received by LOCAL expert before
the grouped GEMMs and undoes that afterwards (``fused_moe/experts/deep_gemm_moe.py:279-329``):

* ``deepgemm_moe_permute`` -> ``ep_scatter`` (``fused_moe/deep_gemm_utils.py:347-420,56-249``): expert ``e``
  owns rows ``[start[e], start[e] + count[e])`` of the output with ``start`` the exclusive prefix sum of the
  counts rounded up to the GEMM's M alignment (128; ``compute_aligned_M :26-44``); every received row is
  copied, with its scales, to one output row per local expert in its top-k; ``expert_ids`` (``m_indices``)
  holds the expert of every occupied output row and -1 in the padding (``:387-392``); ``inv_perm[m, k]`` is
  the output row of entry (m, k).
* ``deepgemm_unpermute_and_reduce`` -> ``ep_gather`` (``deep_gemm_utils.py:252-344,423-440``):
  ``out[m] = bf16( sum_k w[m,k] * y[inv_perm[m,k]] )`` over the entries with a local expert, fp32
  accumulator, k ascending; the Triton statement ``acc += y.to(f32) * w`` is compiled to ONE fused
  multiply-add per element, which is what ``fma32`` below restates.

Order inside an expert: ``ep_scatter`` hands rows out with an atomic counter, so the reference's order is
whatever the scheduler produced.  Here it is DEFINED (as everywhere on this path): received-row order, i.e.
source rank ascending, then source token ascending.  Any order gives the same GEMM results per row; the
tests compare with the reference as per-expert multisets and through ``inv_perm``.
Duplicate expert ids inside one token's top-k are not supported (top-k routing never produces them).
"""

from __future__ import annotations

import numpy as np

from .bits import bf16_to_f32, f32_to_bf16


def expert_starts(per_expert, alignment: int = 128):
    """start[e] (exclusive prefix of the counts rounded up to ``alignment``) and the total M_sum
    (``expert_num_tokens_round_up_and_sum``, deep_gemm_utils.py:17-24)."""
    cnt = np.asarray(per_expert, np.int64)
    padded = (cnt + alignment - 1) // alignment * alignment
    start = np.zeros(cnt.size + 1, np.int64)
    np.cumsum(padded, out=start[1:])
    return start[:-1].copy(), int(start[-1])


def moe_permute(recv_x, recv_scales, recv_topk_idx, num_local_experts: int, alignment: int = 128, m_cap=None):
    """recv_x [M, row] (any dtype, copied verbatim), recv_scales [M, G] float32 or None, recv_topk_idx [M, K]
    LOCAL expert ids or -1.  Returns dict(out_x [m_cap, row], out_scales, expert_ids [m_cap] int32,
    inv_perm [M, K] int32 (-1 where the entry has no local expert), valid [m_cap] bool, m_sum)."""
    x = np.asarray(recv_x)
    ids = np.asarray(recv_topk_idx).astype(np.int64)
    M, K = ids.shape
    ok = (ids >= 0) & (ids < num_local_experts)
    per_expert = np.bincount(ids[ok], minlength=num_local_experts)
    start, m_sum = expert_starts(per_expert, alignment)
    if m_cap is None:
        m_cap = m_sum
    assert m_cap >= m_sum
    out_x = np.zeros((m_cap,) + x.shape[1:], x.dtype)
    out_s = None if recv_scales is None else np.zeros((m_cap, np.asarray(recv_scales).shape[1]), np.float32)
    expert_ids = np.full(m_cap, -1, np.int32)
    inv_perm = np.full((M, K), -1, np.int32)
    valid = np.zeros(m_cap, bool)
    for e in range(num_local_experts):
        rows, ks = np.nonzero(ids == e)           # row-major: received-row order (k is unique per row)
        n = rows.size
        assert n == per_expert[e]
        dst = start[e] + np.arange(n)
        out_x[dst] = x[rows]
        if out_s is not None:
            out_s[dst] = np.asarray(recv_scales, np.float32)[rows]
        expert_ids[dst] = e
        inv_perm[rows, ks] = dst
        valid[dst] = True
    return dict(out_x=out_x, out_scales=out_s, expert_ids=expert_ids, inv_perm=inv_perm, valid=valid, m_sum=m_sum,
                expert_start=start, per_expert=per_expert.astype(np.int32))


def fma32(a, b, c):
    """round_to_float32(a * b + c) with ONE rounding, element-wise (what ``fma.rn.f32`` computes).
    The product of two float32 is exact in float64; the float64 sum is rounded once more when it is narrowed,
    so the rare double-rounding case (the float64 sum lands exactly on a float32 tie while the exact sum does
    not) is repaired with the exact error term of the addition (TwoSum)."""
    a = np.asarray(a, np.float32).astype(np.float64)
    b = np.asarray(b, np.float32).astype(np.float64)
    c = np.asarray(c, np.float32).astype(np.float64)
    a, b, c = np.broadcast_arrays(a, b, c)
    p = a * b                       # exact (24 x 24 bits)
    s = p + c                       # one float64 rounding
    bb = s - p
    err = (p - (s - bb)) + (c - bb)  # exact: p + c == s + err
    r = s.astype(np.float32)
    bits = np.ascontiguousarray(s).view(np.uint64)
    low = bits & np.uint64((1 << 29) - 1)
    tie = (low == np.uint64(1 << 28)) & (err != 0) & np.isfinite(s)
    if tie.any():
        mag = np.abs(s)
        assert (mag[tie] >= 2.0**-126).all(), "fma32: float32-subnormal results are not handled"
        trunc = (bits & ~np.uint64((1 << 29) - 1))
        lo = trunc.view(np.float64)                                   # toward zero
        hi = (trunc + np.uint64(1 << 29)).view(np.float64)            # away from zero
        grows = (err > 0) == (s > 0)                                  # exact magnitude is above |s|
        fixed = np.where(grows, hi, lo).astype(np.float32)
        r = np.where(tie, fixed, r)
    return r.astype(np.float32)


def moe_unpermute_reduce(y_grouped, recv_topk_idx, recv_topk_weights, inv_perm, num_local_experts: int):
    """y_grouped [m_cap, H] bf16 bits (expert outputs in the grouped layout) -> z [M, H] bf16 bits, the row each
    received token sends back: fp32 accumulator, one fused multiply-add per entry, k ascending, one bf16 rounding
    (``ep_gather``, deep_gemm_utils.py:252-296)."""
    y = np.asarray(y_grouped)
    ids = np.asarray(recv_topk_idx).astype(np.int64)
    w = np.asarray(recv_topk_weights, np.float32)
    inv = np.asarray(inv_perm).astype(np.int64)
    M, K = ids.shape
    acc = np.zeros((M, y.shape[1]), np.float32)
    for k in range(K):
        m = (ids[:, k] >= 0) & (ids[:, k] < num_local_experts)
        if not m.any():
            continue
        rows = bf16_to_f32(y[inv[m, k]])
        acc[m] = fma32(rows, w[m, k][:, None], acc[m])
    return f32_to_bf16(acc)
