"""Bit-exact parity checks of the PRODUCTION multi-process calls against the CPU oracle.

Shared by ``tests/mp_worker.py`` (torchrun, one rank per GPU over CUDA IPC) and by ``bench.py``
(which runs one untimed step of exactly the calls it times through these checks before the timed
region, at every N, and refuses to print a number if they fail).  Every rank

  1. builds its own seeded inputs and quantises them with the oracle (``oracle/fp8.py``),
  2. gathers all ranks' oracle inputs over the host group (gloo) and evaluates the oracle for the
     rows IT must receive (``oracle/dispatch.py``),
  3. makes the library call with all phases in one go -- the calls vLLM makes through the
     ``deep_ep`` shim: ``Buffer.dispatch`` / ``Buffer.combine`` (overlapped combine kernel at
     world > 1), the host-sync-free pair, ``ll_dispatch_launch(LL_PHASE_ALL)`` +
     ``low_latency_combine(phases=PHASE_ALL)`` (optionally captured in a CUDA graph and replayed),
  4. compares bytes, scale bits, ids, weight bits, counts and the combined output.

Expert stand-in: rank d answers row r with ``bf16(dequant(r) * (1 + 0.25 d))`` (HT) or
``bf16(dequant(r) * (1 + 0.125 d + 0.5 e_local))`` (LL), computed in numpy from the ORACLE's
received rows, so a row delivered to the wrong place or summed in the wrong order shows up.
"""
from __future__ import annotations

import numpy as np
import torch

from _util import bf16_bits, bits_to_bf16, make_rank_inputs, u8
from oracle import bits, combine, dispatch, fp8, layout, permute


class ParityError(AssertionError):
    pass


_errs: list[str] = []  # mismatches of the check in progress (raised only after all ranks agree, see _agree)


def _eq(a, b, what, rank):
    a, b = np.asarray(a), np.asarray(b)
    if a.shape != b.shape or not np.array_equal(a, b):
        n = int((a != b).sum()) if a.shape == b.shape else -1
        _errs.append(f"rank {rank}: {what} differs from the oracle ({n} of {a.size} elements, shapes {a.shape} vs {b.shape})")
        return False
    return True


def _agree(group, W, me):
    """Collective: raise ParityError on EVERY rank if any rank recorded a mismatch (a rank that
    stopped alone would leave its peers waiting inside the next collective kernel)."""
    mine = list(_errs)
    _errs.clear()
    all_errs = _gather(group, mine, W)
    flat = [e for per_rank in all_errs for e in per_rank]
    if flat:
        raise ParityError("; ".join(flat[:6]) + (f" (+{len(flat) - 6} more)" if len(flat) > 6 else ""))


def _gather(group, obj, W):
    if W == 1:
        return [obj]
    import torch.distributed as dist

    out = [None] * W
    dist.all_gather_object(out, obj, group=group)
    return out


def _ht_scale(d: int) -> np.float32:
    return np.float32(1.0 + 0.25 * d)


def check_ht(buf, group, W, me, dev, Ts, H, E, K, seed, dist="uniform", *, nosync=False, ue8m0=False,
             id_dtype=torch.int64, mask_frac=0.0, combine_out=None):
    """One high-throughput dispatch + combine through the production calls.  Ts[r] = tokens of rank r."""
    x, ids, w = make_rank_inputs(Ts[me], H, E, K, seed * 100 + me, dist=dist, world=W, id_dtype=id_dtype,
                                 mask_frac=mask_frac)
    q, s = fp8.group_quant_fp8(bf16_bits(x), ue8m0=ue8m0, fast=True)
    allr = _gather(group, (q, s, ids.numpy(), w.numpy()), W)
    ref = dispatch.ht_dispatch([a[0] for a in allr], [a[2] for a in allr], [a[3] for a in allr], E,
                               scales=[a[1] for a in allr], dests=[me])[me]
    M = ref["num_recv_tokens"]
    worst = (W * max(Ts)) if nosync else 0
    if nosync and worst == 0:
        return
    rx, rs, ridx, rw, per_expert, h = buf.dispatch(x.to(dev), ids.to(dev), w.to(dev), E, quantize=True, ue8m0=ue8m0,
                                                   num_worst_tokens=worst)
    rsrc = buf.ht_views(h)[4]
    torch.cuda.synchronize()
    tag = f"ht{'_nosync' if nosync else ''}[{dist}, T={Ts[me]}]"
    rows = rx.shape[0]
    if nosync:
        _eq(int(h.num_recv_dev), M, f"{tag} device row count", me)
        _eq(h.recv_per_expert_dev.cpu().numpy(), ref["num_recv_tokens_per_expert"], f"{tag} device per-expert counts", me)
        _eq([len(per_expert), rows], [0, worst], f"{tag} worst-case view sizes", me)
    else:
        _eq(h.num_recv, M, f"{tag} num_recv", me)
        _eq(list(h.rank_prefix), list(ref["rank_prefix"]), f"{tag} rank_prefix", me)
        _eq(list(per_expert), list(ref["num_recv_tokens_per_expert"]), f"{tag} per-expert list", me)
    if rows >= M:
        _eq(u8(rx)[:M], ref["recv_x"], f"{tag} fp8 bytes", me)
        _eq(rs.cpu().numpy()[:M].view(np.uint32), ref["recv_x_scales"].view(np.uint32), f"{tag} scales", me)
        _eq(ridx.cpu().numpy()[:M], ref["recv_topk_idx"], f"{tag} recv_topk_idx", me)
        _eq(rw.cpu().numpy()[:M].view(np.uint32), ref["recv_topk_weights"].view(np.uint32), f"{tag} recv_topk_weights", me)
        _eq(rsrc.cpu().numpy()[:M], ref["recv_src_idx"], f"{tag} recv_src_idx", me)
    # combine (always issued, with whatever row count the library reported, so that the collective completes)
    y_np = bits.f32_to_bf16(fp8.dequant_fp8(ref["recv_x"], ref["recv_x_scales"]) * _ht_scale(me))
    # rows beyond the M received ones (host-sync-free mode) must never be read: poison them
    y = torch.full((rows, H), float("nan"), dtype=torch.bfloat16, device=dev)
    if min(M, rows):
        y[: min(M, rows)] = bits_to_bf16(y_np, dev)[: min(M, rows)]
    out = combine_out if combine_out is not None else torch.empty((Ts[me], H), dtype=torch.bfloat16, device=dev)
    out.fill_(float("nan"))
    buf.combine(y, h, out)
    torch.cuda.synchronize()
    in_rank = layout.dispatch_layout(ids.numpy(), E, W)[2]
    deq = fp8.dequant_fp8(q, s)
    parts = []
    for d in range(W):
        tok = np.nonzero(in_rank[:, d])[0]
        parts.append((tok, bits.f32_to_bf16(deq[tok] * _ht_scale(d))))
    want = combine.ht_combine_source(Ts[me], H, parts)
    _eq(bf16_bits(out), want, f"{tag} combine output", me)
    _agree(group, W, me)


def check_ht_grouped(buf, group, W, me, dev, Ts, H, E, K, seed, dist="uniform", mask_frac=0.0):
    """Dispatch, regroup the received rows by local expert (b200ep_ht_permute) and combine straight from the grouped
    layout (b200ep_ht_combine_permuted, all phases): the calls a DeepGEMM expert class makes through Route B.
    Expert stand-in: the grouped row of entry (t, k) on rank d, local expert e, answers bf16(dequant(row) * (1 + 0.125 d + 0.5 e))."""
    el = E // W
    x, ids, w = make_rank_inputs(Ts[me], H, E, K, seed * 100 + me, dist=dist, world=W, mask_frac=mask_frac)
    q, s = fp8.group_quant_fp8(bf16_bits(x), fast=True)
    allr = _gather(group, (q, s, ids.numpy(), w.numpy()), W)
    ref = dispatch.ht_dispatch([a[0] for a in allr], [a[2] for a in allr], [a[3] for a in allr], E,
                               scales=[a[1] for a in allr], dests=[me])[me]
    p = permute.moe_permute(ref["recv_x"], ref["recv_x_scales"], ref["recv_topk_idx"], el, alignment=128)
    rx, rs, ridx, rw, per_expert, h = buf.dispatch(x.to(dev), ids.to(dev), w.to(dev), E, quantize=True)
    a_perm, s_perm, eids, inv = buf.ht_permute(h, per_expert=per_expert)
    torch.cuda.synchronize()
    tag = f"ht_grouped[{dist}, T={Ts[me]}]"
    ok_shape = _eq([int(a_perm.shape[0]), int(inv.shape[0])], [p["m_sum"], ref["num_recv_tokens"]], f"{tag} sizes", me)
    if ok_shape:
        v = p["valid"]
        _eq(u8(a_perm)[v], p["out_x"][v], f"{tag} grouped fp8 bytes", me)
        _eq(s_perm.cpu().numpy().view(np.uint32)[v], p["out_scales"].view(np.uint32)[v], f"{tag} grouped scales", me)
        _eq(eids.cpu().numpy(), p["expert_ids"], f"{tag} expert_ids", me)
        _eq(inv.cpu().numpy(), p["inv_perm"], f"{tag} inv_perm", me)
    # expert outputs in the grouped layout, from the ORACLE's layout (padding rows are poison: never read)
    yg = np.full((int(a_perm.shape[0]), H), 0x7FC0, np.uint16)
    if ok_shape and p["m_sum"]:
        v = p["valid"]
        f = (1.0 + 0.125 * me + 0.5 * p["expert_ids"][v]).astype(np.float32)
        yg[v] = bits.f32_to_bf16(fp8.dequant_fp8(p["out_x"][v], p["out_scales"][v]) * f[:, None])
    out = torch.full((Ts[me], H), float("nan"), dtype=torch.bfloat16, device=dev)
    inv_dev = inv if ok_shape else torch.full((h.num_recv, K), -1, dtype=torch.int32, device=dev)
    buf.combine_permuted(bits_to_bf16(yg, dev), inv_dev, h, out)
    torch.cuda.synchronize()
    # what MY tokens get back: per destination rank d, z = bf16(sum_k fma(row(t,k), w[t,k])) over d's experts, k ascending
    idn, wn = ids.numpy().astype(np.int64), w.numpy()
    deq = fp8.dequant_fp8(q, s)
    in_rank = layout.dispatch_layout(ids.numpy(), E, W)[2]
    parts = []
    for d in range(W):
        tok = np.nonzero(in_rank[:, d])[0]
        acc = np.zeros((tok.size, H), np.float32)
        for k in range(K):
            e = idn[tok, k]
            m = (e >= 0) & (e < E) & (e // el == d)
            if m.any():
                f = (1.0 + 0.125 * d + 0.5 * (e[m] % el)).astype(np.float32)
                rows = bits.bf16_to_f32(bits.f32_to_bf16(deq[tok[m]] * f[:, None]))
                acc[m] = permute.fma32(rows, wn[tok[m], k][:, None], acc[m])
        parts.append((tok, bits.f32_to_bf16(acc)))
    _eq(bf16_bits(out), combine.ht_combine_source(Ts[me], H, parts), f"{tag} combine output", me)
    _agree(group, W, me)


def _ll_scale(d: int, el: int) -> np.float32:
    return np.float32(1.0 + 0.125 * d + 0.5 * el)


class LlGraphCase:
    """Low-latency dispatch + combine with all phases per call, optionally captured ONCE in a CUDA
    graph and replayed with fresh inputs each time (what ``bench.py`` times)."""

    def __init__(self, buf, group, W, me, dev, T, H, E, K, max_tok, *, use_fp8=True, ue8m0=False, graph=True, slot=0):
        from b200ep import LlHandle
        from b200ep import _lib as L

        self.buf, self.group, self.W, self.me, self.dev = buf, group, W, me, dev
        self.T, self.H, self.E, self.K, self.max_tok, self.use_fp8, self.ue8m0 = T, H, E, K, max_tok, use_fp8, ue8m0
        self.el, self.cap = E // W, W * max_tok
        self.x = torch.zeros(T, H, dtype=torch.bfloat16, device=dev)
        self.ids = torch.zeros(T, K, dtype=torch.int64, device=dev)
        self.w = torch.zeros(T, K, dtype=torch.float32, device=dev)
        self.y = torch.zeros(self.el, self.cap, H, dtype=torch.bfloat16, device=dev)
        self.out = torch.zeros(T, H, dtype=torch.bfloat16, device=dev)
        self.h = LlHandle(slot, T, H, K, E, max_tok, use_fp8)
        self.L = L
        self.graph = None
        self.use_graph = graph
        self.stream = torch.cuda.Stream(device=dev)

    def _launch(self):
        self.buf.ll_dispatch_launch(self.h, self.x, self.ids, self.ue8m0, self.L.LL_PHASE_ALL)
        self.buf.low_latency_combine(self.y, self.ids, self.w, self.h, self.out, phases=self.L.PHASE_ALL)

    def run(self, seed, dist="uniform", mask_frac=0.0, barrier=None):
        W, me, T, H, E, K, el, cap = self.W, self.me, self.T, self.H, self.E, self.K, self.el, self.cap
        x, ids, w = make_rank_inputs(T, H, E, K, seed * 100 + me, dist=dist, world=W, mask_frac=mask_frac)
        xb = bf16_bits(x)
        if self.use_fp8:
            q, s = fp8.group_quant_fp8(xb, ue8m0=self.ue8m0, fast=True)
        else:
            q, s = xb, None
        allr = _gather(self.group, (q, s, ids.numpy()), W)
        ref = dispatch.ll_dispatch([a[0] for a in allr], [a[2] for a in allr], E, self.max_tok,
                                   scales=[a[1] for a in allr] if self.use_fp8 else None, dests=[me])[me]

        def deq(rows_q, rows_s):
            return fp8.dequant_fp8(rows_q, rows_s) if self.use_fp8 else bits.bf16_to_f32(rows_q)

        y = np.zeros((el, cap, H), np.uint16)
        for e in range(el):
            n = int(ref["recv_count"][e])
            if n:
                sc = ref["packed_recv_scales"][e, :n] if self.use_fp8 else None
                y[e, :n] = bits.f32_to_bf16(deq(ref["packed_recv_x"][e, :n], sc) * _ll_scale(me, e))
        self.x.copy_(x); self.ids.copy_(ids); self.w.copy_(w)
        self.y.copy_(bits_to_bf16(y.reshape(-1, H), self.dev).view(el, cap, H))
        self.out.fill_(float("nan"))
        torch.cuda.synchronize()
        if barrier is not None:
            barrier()
        with torch.cuda.stream(self.stream):
            if self.use_graph and self.graph is None:
                self._launch()  # the first call on a context sets function attributes: keep it out of the capture
                self.stream.synchronize()
                if barrier is not None:
                    barrier()
                self.out.fill_(float("nan"))
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=self.stream, capture_error_mode="relaxed"):
                    self._launch()
                self.graph = g
                if barrier is not None:
                    barrier()
            if self.use_graph:
                self.graph.replay()
            else:
                self._launch()
            self.stream.synchronize()
        px, ps, cnt, src, rng = self.buf.ll_views(self.h)
        tag = f"ll[{dist}, T={T}, E={E}{', graph' if self.use_graph else ''}]"
        _eq(cnt.cpu().numpy(), ref["recv_count"], f"{tag} recv_count", me)
        _eq(rng.cpu().numpy(), ref["layout_range"], f"{tag} layout_range", me)
        pxh = u8(px) if self.use_fp8 else bf16_bits(px)
        psh = ps.cpu().numpy().view(np.uint32) if self.use_fp8 else None
        srch = src.cpu().numpy()
        for e in range(el):
            n = int(ref["recv_count"][e])
            _eq(pxh[e, :n], ref["packed_recv_x"][e, :n], f"{tag} payload of local expert {e}", me)
            _eq(srch[e, :n], ref["src_info"][e, :n], f"{tag} src_info of local expert {e}", me)
            if self.use_fp8:
                _eq(psh[e, :n], ref["packed_recv_scales"][e, :n].view(np.uint32), f"{tag} scales of local expert {e}", me)
        # expected combine for MY tokens: entry (t, k) comes back as bf16(deq(row t) * scale(owner, e_local))
        mine = deq(q, s)
        idn = ids.numpy().astype(np.int64)
        rows = np.zeros((T, K, H), np.uint16)
        for k in range(K):
            ok = (idn[:, k] >= 0) & (idn[:, k] < E)
            f = np.ones(T, np.float32)
            f[ok] = (1.0 + 0.125 * (idn[ok, k] // el) + 0.5 * (idn[ok, k] % el)).astype(np.float32)
            rows[:, k] = bits.f32_to_bf16(mine * f[:, None])
        want = combine.ll_combine_source(idn, w.numpy(), rows, E)
        _eq(bf16_bits(self.out), want, f"{tag} combine output", me)
        _agree(self.group, W, me)


def run_bench_parity(buf, group, W, me, dev, barrier, *, H=7168, E=256, K=8, T_ht=1100, T_ll=128, ll_experts=(256,)):
    """The block ``bench.py`` runs before its timed region.  Returns the dict for the JSON line.
    Raises ParityError on every rank (the ranks agree on the verdict after each check) on any mismatch."""
    res = {"shape": {"hidden": H, "experts": E, "topk": K, "ht_tokens_per_rank": T_ht, "ll_tokens_per_rank": T_ll,
                     "ep_size": W},
           "how": "one untimed step through exactly the timed calls, compared bit for bit with oracle/ on every rank"}
    Ts = [T_ht - 13 * r for r in range(W)]
    for i, dist_kind in enumerate(("uniform", "zipf")):
        check_ht(buf, group, W, me, dev, Ts, H, E, K, seed=700 + i, dist=dist_kind)
        check_ht(buf, group, W, me, dev, Ts, H, E, K, seed=700 + i, dist=dist_kind)  # second slot, next epoch
    res["ht"] = "ok"
    for i, dist_kind in enumerate(("uniform", "one_rank")):
        check_ht(buf, group, W, me, dev, Ts, H, E, K, seed=720 + i, dist=dist_kind, nosync=True)
    res["ht_nosync"] = "ok"
    # the expert-grouped calls are not part of the timed step: a mismatch there is reported (and bench.py then skips their
    # timing) but does not withhold the step's number.  Every rank reaches the same verdict (_agree is collective).
    try:
        for i, dist_kind in enumerate(("uniform", "zipf")):
            check_ht_grouped(buf, group, W, me, dev, Ts, H, E, K, seed=730 + i, dist=dist_kind, mask_frac=0.1 * i)
        res["ht_grouped"] = "ok"
    except ParityError as ex:
        res["ht_grouped"] = "FAILED: " + str(ex)[:500]
    for ne in ll_experts:
        if ne % W:
            continue
        case = LlGraphCase(buf, group, W, me, dev, T_ll, H, ne, K, T_ll, graph=True)
        for i, dist_kind in enumerate(("uniform", "zipf", "uniform")):
            case.run(seed=740 + i, dist=dist_kind, mask_frac=0.1 * i, barrier=barrier)
    res["ll"] = "ok"
    return res
