"""Token dispatch, CPU restatement for W ranks in one process (test infrastructure).

High-throughput contract (what vLLM reads back from ``deep_ep.Buffer.dispatch``,
``prepare_finalize/deepep_ht.py:139-162,183-253``):
  recv_x / recv_x_scales  one copy of a token per destination rank that owns at
                          least one of its top-k experts,
  recv_topk_idx           ids rewritten to the destination's LOCAL expert ids,
                          -1 for experts that live elsewhere (deepep_ht.py:204-220
                          maps them back to global ids),
  recv_topk_weights       the router weights, 0 where recv_topk_idx is -1,
  num_recv_tokens_per_expert_list  per local expert (deepep_ht.py:225-227).
The data movement is the gather of ``cpu_fused_moe_torch``
(``cpu_fused_moe.py:447-449``: ``sorted_tokens = input[idxs // K]``) split by
owner rank, i.e. the all-gather of ``AgRsAll2AllManager.dispatch``
(``distributed/device_communicators/all2all.py:84-122``) restricted to the rows
a rank needs.

Low-latency contract (``deepep_ll.py:292-342``; layout = BatchedExperts as built
by ``prepare_finalize/batched.py:82-152``): ``packed_recv_x[e_local, slot, :]``
with the tokens of an expert packed at slots ``[0, count)``.

Canonical order -- DEFINED here because no runnable reference
defines one: source rank ascending, then source token ascending.  At W=1 this is
exactly the boolean-mask gather of ``batched.py:115-122``.
"""

from __future__ import annotations

import numpy as np

from .layout import dispatch_layout, expert_owner


def ht_dispatch(xs, topk_ids, topk_weights, num_experts: int, scales=None, dests=None):
    """xs[s]: [T_s, row] payload rows of source rank s (any dtype; copied verbatim).
    scales[s]: optional [T_s, G] float32.  topk_ids[s]: [T_s, K]; topk_weights[s]:
    [T_s, K] float32.  Returns one dict per destination rank (``None`` for ranks not in
    ``dests`` when that is given: a rank of a multi-process test only needs its own)."""
    W = len(xs)
    E_local = num_experts // W
    lay = [dispatch_layout(topk_ids[s], num_experts, W) for s in range(W)]
    out = []
    for d in range(W):
        if dests is not None and d not in dests:
            out.append(None)
            continue
        rx, rs, ridx, rw, rsrc = [], [], [], [], []
        prefix = np.zeros(W + 1, np.int64)
        for s in range(W):
            in_rank = lay[s][2][:, d].astype(bool)
            rows = np.nonzero(in_rank)[0]
            prefix[s + 1] = prefix[s] + rows.size
            ids = np.asarray(topk_ids[s]).astype(np.int64)[rows]
            own = expert_owner(ids, num_experts, W) == d
            rx.append(np.asarray(xs[s])[rows])
            if scales is not None:
                rs.append(np.asarray(scales[s])[rows])
            ridx.append(np.where(own, ids - d * E_local, -1).astype(np.int64))
            rw.append(np.where(own, np.asarray(topk_weights[s], np.float32)[rows], np.float32(0)))
            rsrc.append(rows.astype(np.int32))
        recv_idx = np.concatenate(ridx, axis=0)
        per_expert = np.bincount(recv_idx[recv_idx >= 0], minlength=E_local).astype(np.int32)
        out.append(
            dict(
                recv_x=np.concatenate(rx, axis=0),
                recv_x_scales=np.concatenate(rs, axis=0) if scales is not None else None,
                recv_topk_idx=recv_idx,
                recv_topk_weights=np.concatenate(rw, axis=0).astype(np.float32),
                recv_src_idx=np.concatenate(rsrc, axis=0),
                rank_prefix=prefix,  # rows [prefix[s], prefix[s+1]) came from rank s
                num_recv_tokens=int(prefix[W]),
                num_recv_tokens_per_expert=per_expert,
            )
        )
    return out


def ll_dispatch(xs, topk_ids, num_experts: int, max_tokens_per_rank: int, scales=None, dests=None):
    """Low-latency dispatch.  Returns one dict per destination rank with
    packed_recv_x [E_local, W*max_tok, row], optional packed scales, recv_count
    [E_local], src_info [E_local, W*max_tok] (= t*K + k of the source entry, -1 if
    unused) and layout_range [E_local, W, 2] (start, count per source rank).
    ``dests`` restricts the work to some destination ranks (``None`` elsewhere)."""
    W = len(xs)
    E_local = num_experts // W
    cap = W * max_tokens_per_rank
    out = []
    for d in range(W):
        if dests is not None and d not in dests:
            out.append(None)
            continue
        row = np.asarray(xs[0]).shape[1]
        px = np.zeros((E_local, cap, row), np.asarray(xs[0]).dtype)
        ps = None
        if scales is not None:
            ps = np.zeros((E_local, cap, np.asarray(scales[0]).shape[1]), np.float32)
        cnt = np.zeros(E_local, np.int32)
        src_info = np.full((E_local, cap), -1, np.int32)
        rng = np.zeros((E_local, W, 2), np.int32)
        for el in range(E_local):
            e = d * E_local + el
            for s in range(W):
                ids = np.asarray(topk_ids[s]).astype(np.int64)
                T, K = ids.shape
                assert T <= max_tokens_per_rank
                hit = ids == e
                rows = np.nonzero(hit.any(axis=1))[0]
                ks = hit[rows].argmax(axis=1)
                n = rows.size
                a = int(cnt[el])
                rng[el, s] = (a, n)
                px[el, a : a + n] = np.asarray(xs[s])[rows]
                if ps is not None:
                    ps[el, a : a + n] = np.asarray(scales[s])[rows]
                src_info[el, a : a + n] = rows * K + ks
                cnt[el] = a + n
        out.append(
            dict(packed_recv_x=px, packed_recv_scales=ps, recv_count=cnt, src_info=src_info, layout_range=rng)
        )
    return out
