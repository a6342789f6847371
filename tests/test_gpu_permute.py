"""GPU parity of the expert-grouped layout, through the C ABI:
b200ep_ht_permute against oracle/permute.py (bit-exact) and against vLLM's own ep_scatter (same rows per expert),
b200ep_ht_combine_permuted against the oracle (bit-exact) and against vLLM's ep_gather + the plain combine."""
import numpy as np
import pytest
import torch

from _util import bf16_bits, bits_to_bf16, make_rank_inputs, u8
from oracle import bits, combine, dispatch, fp8, permute

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _inputs(W, Ts, H, E, K, seed, dist="uniform", mask=0.0):
    xs, ids, ws = [], [], []
    for r in range(W):
        x, i, w = make_rank_inputs(Ts[r], H, E, K, seed * 100 + r, dist=dist, world=W, mask_frac=mask)
        xs.append(x); ids.append(i); ws.append(w)
    return xs, ids, ws


def _expert_outputs(p, H, d):
    """Deterministic stand-in for the grouped GEMM output: bf16 rows that depend on the grouped input row."""
    rng = np.random.default_rng(1234 + d)
    y = rng.standard_normal((p["out_x"].shape[0], H)).astype(np.float32)
    y += (p["out_x"][:, :1].astype(np.float32) - 128.0) / 64.0
    y[~p["valid"]] = np.nan  # padding rows must never be read
    return bits.f32_to_bf16(y)


@pytest.mark.parametrize("W", [1, 2, 4, 8])
@pytest.mark.parametrize("kind", ["quant", "bf16"])
def test_permute_and_permuted_combine_bit_exact(W, kind):
    from b200ep import LocalWorld

    H, E, K = 1024, 32, 8
    world = LocalWorld(W, sym_bytes=192 << 20)
    try:
        for it, (dist, mask) in enumerate([("uniform", 0.0), ("zipf", 0.2), ("one_rank", 0.0)]):
            Ts = [max(0, 301 - 37 * r) if r != 1 else 0 for r in range(W)] if W > 1 else [257]
            xs, ids, ws = _inputs(W, Ts, H, E, K, seed=it + 10 * W, dist=dist, mask=mask)
            xb = [bf16_bits(x) for x in xs]
            idn, wn = [i.numpy() for i in ids], [w.numpy() for w in ws]
            if kind == "bf16":
                ref = dispatch.ht_dispatch(xb, idn, wn, E)
                res = world.dispatch([x.to(DEV) for x in xs], [i.to(DEV) for i in ids], [w.to(DEV) for w in ws], E)
            else:
                qs = [fp8.group_quant_fp8(b, formula="div") for b in xb]
                ref = dispatch.ht_dispatch([q for q, _ in qs], idn, wn, E, scales=[s for _, s in qs])
                res = world.dispatch([x.to(DEV) for x in xs], [i.to(DEV) for i in ids], [w.to(DEV) for w in ws], E,
                                     quantize=True)
            el = E // W
            perms, ys, invs, zs = [], [], [], []
            for d in range(W):
                o = ref[d]
                p = permute.moe_permute(o["recv_x"], o["recv_x_scales"], o["recv_topk_idx"], el, alignment=128)
                h = res[d][-1]
                ax, asc, eids, inv = world.bufs[d].ht_permute(h, per_expert=res[d][5])
                assert ax.shape[0] == p["m_sum"]
                v = p["valid"]
                got_x = bf16_bits(ax) if kind == "bf16" else u8(ax)
                assert np.array_equal(got_x[v], p["out_x"][v]), "grouped rows differ from the oracle"
                if kind != "bf16":
                    assert np.array_equal(asc.cpu().numpy().view(np.uint32)[v], p["out_scales"].view(np.uint32)[v])
                assert np.array_equal(eids.cpu().numpy(), p["expert_ids"])
                assert np.array_equal(inv.cpu().numpy(), p["inv_perm"])
                y = _expert_outputs(p, H, d)
                perms.append(p); ys.append(y); invs.append(inv)
                zs.append(permute.moe_unpermute_reduce(y, o["recv_topk_idx"], o["recv_topk_weights"], p["inv_perm"], el))
            outs = world.combine_permuted([bits_to_bf16(y, DEV) for y in ys], invs, [r[-1] for r in res])
            ref_out = combine.ht_combine(zs, ref, Ts)
            for r in range(W):
                assert np.array_equal(bf16_bits(outs[r]), ref_out[r]), f"permuted combine differs on rank {r}"
            # the worst-case sizing of a host-sync-free caller gives the same layout in larger buffers
            d = W - 1
            ax2, _, eids2, inv2 = world.bufs[d].ht_permute(res[d][-1])
            assert ax2.shape[0] >= perms[d]["m_sum"]
            assert np.array_equal(inv2.cpu().numpy(), perms[d]["inv_perm"])
            e2 = eids2.cpu().numpy()
            assert np.array_equal(e2[: perms[d]["m_sum"]], perms[d]["expert_ids"]) and np.all(e2[perms[d]["m_sum"]:] == -1)
    finally:
        world.destroy()


def test_permute_matches_vllm_ep_scatter_and_ep_gather():
    """The reference functions themselves (vLLM fused_moe/deep_gemm_utils.py ep_scatter / ep_gather, Triton) on the same
    received rows: same rows per expert (the reference's order inside an expert is arbitrary), same expert_ids, and the
    same bf16 rows out of the weighted reduction."""
    dgu = pytest.importorskip("vllm.model_executor.layers.fused_moe.deep_gemm_utils")
    from b200ep import LocalWorld

    T, H, E, K, W = 1500, 7168, 256, 8, 1
    world = LocalWorld(W, sym_bytes=1 << 30)
    try:
        x, ids, w = make_rank_inputs(T, H, E, K, 5, mask_frac=0.1)
        buf = world.bufs[0]
        rx, rs, ridx, rw, per_expert, h = buf.dispatch(x.to(DEV), ids.to(DEV), w.to(DEV), E, quantize=True)
        ax, asc, eids, inv = buf.ht_permute(h, per_expert=per_expert)
        m_sum = ax.shape[0]
        cnt = torch.tensor(per_expert, dtype=torch.int32, device=DEV)
        start = torch.empty(E, dtype=torch.int32, device=DEV)
        vx = torch.zeros((m_sum, H), dtype=torch.float8_e4m3fn, device=DEV)
        vs = torch.zeros((m_sum, H // 128), dtype=torch.float32, device=DEV)
        veid = torch.full((m_sum,), -1, dtype=torch.int32, device=DEV)
        vinv = torch.full((h.num_recv, K), -1, dtype=torch.int32, device=DEV)
        dgu.ep_scatter(recv_x=rx, recv_x_scale=rs, recv_topk=ridx, num_recv_tokens_per_expert=cnt, expert_map=None,
                       expert_start_loc=start, output_tensor=vx, output_tensor_scale=vs, m_indices=veid, output_index=vinv)
        torch.cuda.synchronize()
        assert torch.equal(veid, eids)
        a, b = u8(ax), u8(vx)
        sa, sb = asc.cpu().numpy().view(np.uint32), vs.cpu().numpy().view(np.uint32)
        e_np = eids.cpu().numpy()
        for e in range(E):
            rows = np.nonzero(e_np == e)[0]
            if rows.size == 0:
                continue
            ka = np.concatenate([a[rows], sa[rows].view(np.uint8)], 1)
            kb = np.concatenate([b[rows], sb[rows].view(np.uint8)], 1)
            ka = ka[np.lexsort(ka.T[::-1])]
            kb = kb[np.lexsort(kb.T[::-1])]
            assert np.array_equal(ka, kb), f"expert {e}: different rows than ep_scatter"
        # weighted reduction: our kernel on our layout vs ep_gather on the reference's layout, same expert outputs per entry
        g = torch.Generator(device="cpu").manual_seed(3)
        y_entry = torch.randn(h.num_recv, K, H, generator=g).to(torch.bfloat16).to(DEV)   # output of entry (m, k)
        ours = torch.zeros((m_sum, H), dtype=torch.bfloat16, device=DEV)
        theirs = torch.zeros((m_sum, H), dtype=torch.bfloat16, device=DEV)
        ok = ridx >= 0
        ours[inv[ok].long()] = y_entry[ok]
        theirs[vinv[ok].long()] = y_entry[ok]
        z_ref = torch.empty((h.num_recv, H), dtype=torch.bfloat16, device=DEV)
        dgu.ep_gather(input_tensor=theirs, recv_topk_ids=ridx, recv_topk_weight=rw, input_index=vinv, expert_map=None,
                      output_tensor=z_ref)
        out_ref = buf.combine(z_ref, h)
        out = buf.combine_permuted(ours, inv, h)
        torch.cuda.synchronize()
        assert torch.equal(out.view(torch.int16), out_ref.view(torch.int16)), "differs from ep_gather + combine"
    finally:
        world.destroy()


def test_permute_full_size_world8_shape():
    """BASELINE shape at EP=8 on the receiving side: 32 local experts, ~21.7k received rows of 7168 e4m3 bytes.
    Emulated with one rank receiving from itself only would not reach that row count, so W=1 with 32 experts and
    4096 x top-8 tokens (32768 grouped rows) stands in: counts, alignment and a round trip through the grouped layout."""
    from b200ep import LocalWorld

    T, H, E, K = 4096, 7168, 32, 8
    world = LocalWorld(1, sym_bytes=1 << 30)
    try:
        x, ids, w = make_rank_inputs(T, H, E, K, 11)
        buf = world.bufs[0]
        rx, rs, ridx, rw, per_expert, h = buf.dispatch(x.to(DEV), ids.to(DEV), w.to(DEV), E, quantize=True)
        ax, asc, eids, inv = buf.ht_permute(h, per_expert=per_expert)
        assert ax.shape[0] == sum((n + 127) // 128 * 128 for n in per_expert)
        ok = ridx >= 0
        assert int(ok.sum()) == T * K
        # gather back through inv_perm: every entry's grouped row equals its received row
        m_of = torch.arange(h.num_recv, device=DEV)[:, None].expand(-1, K)[ok]
        assert torch.equal(ax.view(torch.uint8)[inv[ok].long()], rx.view(torch.uint8)[m_of])
        assert torch.equal(asc[inv[ok].long()], rs[m_of])
        assert torch.equal(eids[inv[ok].long()].long(), ridx[ok])
        # identity experts (dequantised rows), weights summing to one: the permuted combine returns ~x
        deq = (ax.float().view(-1, H // 128, 128) * asc[:, :, None]).view(-1, H).to(torch.bfloat16)
        out = buf.combine_permuted(deq, inv, h)
        ref = x.to(DEV).float()
        err = (out.float() - ref).abs().max().item()
        assert err <= 0.08 * ref.abs().max().item(), err
    finally:
        world.destroy()


def test_route_b_drop_ins_against_vllm_functions(monkeypatch):
    """Route B: b200ep.vllm_backend.deepgemm_moe_permute / deepgemm_unpermute_and_reduce called exactly as
    DeepGemmExperts.apply calls vLLM's functions of the same names (deep_gemm_moe.py:279-291,322-329), on rows
    received through the deep_ep shim; the deferred unpermute is picked up by deep_ep.Buffer.combine.  Reference:
    vLLM's own two functions + the plain combine on the same dispatch."""
    dgu = pytest.importorskip("vllm.model_executor.layers.fused_moe.deep_gemm_utils")
    import torch.distributed as dist
    import vllm.model_executor.layers.fused_moe.modular_kernel as mk

    import deep_ep
    from b200ep import vllm_backend as vb

    monkeypatch.setattr(dgu, "get_mk_alignment_for_contiguous_layout", lambda: [128, 128])  # DeepGEMM is not installed here
    if not dist.is_initialized():
        dist.init_process_group("gloo", rank=0, world_size=1, init_method="tcp://127.0.0.1:29547")
    torch.cuda.set_device(0)
    T, H, E, K = 700, 2048, 64, 8
    x, ids, w = make_rank_inputs(T, H, E, K, seed=21, mask_frac=0.1)
    buf = deep_ep.Buffer(group=dist.group.WORLD, num_nvl_bytes=1 << 30, num_rdma_bytes=0, low_latency_mode=False,
                         num_qps_per_rank=1, explicitly_destroy=True)
    try:
        tpr, _, tpe, inr, _ = buf.get_dispatch_layout(topk_idx=ids.to(DEV), num_experts=E)
        (rx, rs), ridx, rw, per_expert, handle, _ = buf.dispatch(
            x=x.to(DEV), num_tokens_per_rank=tpr, is_token_in_rank=inr, num_tokens_per_expert=tpe, topk_idx=ids.to(DEV),
            topk_weights=w.to(DEV), b200ep_quantize=True)
        meta = mk.ExpertTokensMetadata.make_from_list(per_expert, device=DEV)
        m_sum = dgu.compute_aligned_M(rx.shape[0], K, E, 128, meta)
        ws = torch.empty((m_sum, H), dtype=torch.float8_e4m3fn, device=DEV)
        a1, s1, e1, inv1 = vb.deepgemm_moe_permute(aq=rx, aq_scale=rs, topk_ids=ridx, local_num_experts=E, expert_map=None,
                                                   expert_tokens_meta=meta, aq_out=ws)
        assert a1.data_ptr() == ws.data_ptr() and a1.shape == (m_sum, H)
        a0, s0, e0, inv0 = dgu.deepgemm_moe_permute(aq=rx, aq_scale=rs, topk_ids=ridx, local_num_experts=E, expert_map=None,
                                                    expert_tokens_meta=meta)
        torch.cuda.synchronize()
        assert torch.equal(e0, e1)
        ok = ridx >= 0
        m_of = torch.arange(rx.shape[0], device=DEV)[:, None].expand(-1, K)[ok]
        for a, s, inv in ((a1, s1, inv1), (a0, s0, inv0)):  # both layouts hold every entry's row at inv_perm
            assert torch.equal(a.view(torch.uint8)[inv[ok].long()], rx.view(torch.uint8)[m_of])
            assert torch.equal(s[inv[ok].long()], rs[m_of])
        g = torch.Generator().manual_seed(9)
        y_entry = torch.randn(rx.shape[0], K, H, generator=g).to(torch.bfloat16).to(DEV)
        mm_ours = torch.zeros((m_sum, H), dtype=torch.bfloat16, device=DEV)
        mm_ref = torch.zeros((m_sum, H), dtype=torch.bfloat16, device=DEV)
        mm_ours[inv1[ok].long()] = y_entry[ok]
        mm_ref[inv0[ok].long()] = y_entry[ok]
        out_ref = torch.empty((rx.shape[0], H), dtype=torch.bfloat16, device=DEV)
        dgu.deepgemm_unpermute_and_reduce(a=mm_ref, topk_ids=ridx, topk_weights=rw, inv_perm=inv0, expert_map=None,
                                          output=out_ref)
        comb_ref, _, _ = buf.combine(x=out_ref, handle=handle)
        comb_ref = comb_ref.clone()
        out_unused = torch.full((rx.shape[0], H), float("nan"), dtype=torch.bfloat16, device=DEV)
        assert vb.deepgemm_unpermute_and_reduce(a=mm_ours, topk_ids=ridx, topk_weights=rw, inv_perm=inv1, expert_map=None,
                                                output=out_unused) is None
        assert handle.deferred is not None
        comb, _, _ = buf.combine(x=out_unused, handle=handle)
        torch.cuda.synchronize()
        assert handle.deferred is None
        assert torch.equal(comb.view(torch.int16), comb_ref.view(torch.int16))
        # rows that did not come from a b200ep dispatch fall through to vLLM's function
        other = rx.clone()
        a2, _, e2, _ = vb.deepgemm_moe_permute(aq=other, aq_scale=rs, topk_ids=ridx, local_num_experts=E, expert_map=None,
                                               expert_tokens_meta=meta)
        assert torch.equal(e2, e0)
    finally:
        buf.destroy()
