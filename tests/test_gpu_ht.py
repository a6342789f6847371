"""GPU parity: high-throughput dispatch + combine vs the CPU oracle, through the C ABI.
W > 1 is emulated on one GPU (b200ep.LocalWorld: phases run rank by rank)."""
import numpy as np
import pytest
import torch

from _util import bf16_bits, bits_to_bf16, make_rank_inputs, u8
from oracle import bits, combine, dispatch, fp8

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _inputs(W, Ts, H, E, K, seed, dist="uniform", idt=torch.int64, mask=0.0):
    xs, ids, ws = [], [], []
    for r in range(W):
        x, i, w = make_rank_inputs(Ts[r], H, E, K, seed * 100 + r, dist=dist, world=W, id_dtype=idt, mask_frac=mask)
        xs.append(x); ids.append(i); ws.append(w)
    return xs, ids, ws


def _check_dispatch(res, ref, kind, K):
    for d, (rx, rs, ridx, rw, rsrc, per_expert, h) in enumerate(res):
        o = ref[d]
        assert h.num_recv == o["num_recv_tokens"]
        assert list(h.rank_prefix) == list(o["rank_prefix"])
        assert per_expert == list(o["num_recv_tokens_per_expert"])
        if kind == "bf16":
            assert np.array_equal(bf16_bits(rx), o["recv_x"])
        else:
            assert np.array_equal(u8(rx), o["recv_x"])
            assert np.array_equal(rs.cpu().numpy().view(np.uint32), o["recv_x_scales"].view(np.uint32))
        assert np.array_equal(ridx.cpu().numpy(), o["recv_topk_idx"])
        assert np.array_equal(rw.cpu().numpy().view(np.uint32), o["recv_topk_weights"].view(np.uint32))
        assert np.array_equal(rsrc.cpu().numpy(), o["recv_src_idx"])


@pytest.mark.parametrize("W", [1, 2, 4, 8])
@pytest.mark.parametrize("kind", ["quant", "fp8", "bf16"])
def test_ht_dispatch_combine_bit_exact(W, kind):
    from b200ep import LocalWorld

    H, E, K = 1024, 32, 8   # BASELINE config #1 shape (granite): 32 experts, top-8, hidden 1024
    world = LocalWorld(W, sym_bytes=192 << 20)
    try:
        for it, (dist, mask, idt) in enumerate([("uniform", 0.0, torch.int64), ("one_rank", 0.0, torch.int32),
                                               ("zipf", 0.2, torch.int64)]):
            Ts = [max(0, 301 - 37 * r) if r != 1 else 0 for r in range(W)]  # ragged, one empty rank
            if W == 1:
                Ts = [257]
            xs, ids, ws = _inputs(W, Ts, H, E, K, seed=it + 10 * W, dist=dist, idt=idt, mask=mask)
            xb = [bf16_bits(x) for x in xs]
            if kind == "bf16":
                ref = dispatch.ht_dispatch(xb, [i.numpy() for i in ids], [w.numpy() for w in ws], E)
                res = world.dispatch([x.to(DEV) for x in xs], [i.to(DEV) for i in ids], [w.to(DEV) for w in ws], E)
            else:
                qs = [fp8.group_quant_fp8(b, formula="div") for b in xb]
                ref = dispatch.ht_dispatch([q for q, _ in qs], [i.numpy() for i in ids], [w.numpy() for w in ws], E,
                                           scales=[s for _, s in qs])
                if kind == "quant":
                    res = world.dispatch([x.to(DEV) for x in xs], [i.to(DEV) for i in ids], [w.to(DEV) for w in ws], E,
                                         quantize=True)
                else:
                    xq = [torch.from_numpy(q).view(torch.float8_e4m3fn).to(DEV) for q, _ in qs]
                    sq = [torch.from_numpy(s).to(DEV) for _, s in qs]
                    res = world.dispatch(xq, [i.to(DEV) for i in ids], [w.to(DEV) for w in ws], E, x_scales=sq)
            _check_dispatch(res, ref, kind, K)
            # expert stand-in: y = bf16(row-dependent function of what was received)
            ys_np = []
            for d in range(W):
                o = ref[d]
                if kind == "bf16":
                    v = bits.bf16_to_f32(o["recv_x"])
                else:
                    v = fp8.dequant_fp8(o["recv_x"], o["recv_x_scales"])
                v = v * (1.0 + 0.25 * d) + 0.125 * o["recv_topk_weights"].sum(1, keepdims=True)
                ys_np.append(bits.f32_to_bf16(v.astype(np.float32)))
            outs = world.combine([bits_to_bf16(y, DEV) for y in ys_np], [r[-1] for r in res])
            ref_out = combine.ht_combine(ys_np, ref, Ts)
            for r in range(W):
                assert np.array_equal(bf16_bits(outs[r]), ref_out[r]), f"combine differs on rank {r}"
    finally:
        world.destroy()


@pytest.mark.parametrize("W", [1, 2])
def test_ht_full_size_properties(W):
    """BASELINE config #2 shape (DeepSeek-V3 layer: hidden 7168, 256 experts, top-8, 4096 tokens/rank).
    Size-independent checks: counts add up, source indices ascending per source rank, every received
    row dequantises to its source row, combine(identity experts) == (#destination ranks) * dequant(x)."""
    from b200ep import LocalWorld

    T, H, E, K = 4096, 7168, 256, 8
    world = LocalWorld(W, sym_bytes=(1 << 30) if W == 1 else (1 << 30))
    try:
        xs, ids, ws = _inputs(W, [T] * W, H, E, K, seed=77)
        res = world.dispatch([x.to(DEV) for x in xs], [i.to(DEV) for i in ids], [w.to(DEV) for w in ws], E, quantize=True)
        owner = [(i // (E // W)) for i in ids]
        ndst = [torch.stack([(o == d).any(1) for d in range(W)], 1).sum(1) for o in owner]
        total = 0
        ys = []
        for d, (rx, rs, ridx, rw, rsrc, per_expert, h) in enumerate(res):
            M = h.num_recv
            total += M
            assert sum(per_expert) == int((ridx >= 0).sum())
            deq = (rx.float().view(M, H // 128, 128) * rs[:, :, None]).view(M, H)
            for s in range(W):
                a, b = h.rank_prefix[s], h.rank_prefix[s + 1]
                t = rsrc[a:b].long()
                assert bool((t[1:] > t[:-1]).all())
                src = xs[s].to(DEV)[t].float()
                # e4m3 with a per-128 scale: relative step 2^-3 of the group amax at worst
                amax = src.view(b - a, H // 128, 128).abs().amax(2, keepdim=True).expand(-1, -1, 128).reshape(b - a, H)
                assert bool(((deq[a:b] - src).abs() <= amax * 0.0625 + 1e-6).all())
            assert bool(((rw > 0) == (ridx >= 0)).all())
            ys.append(deq.to(torch.bfloat16))
        assert total == int(sum(int(n.sum()) for n in ndst))
        outs = world.combine(ys, [r[-1] for r in res])
        for r in range(W):
            ref = xs[r].to(DEV).float() * ndst[r].to(DEV)[:, None]
            err = (outs[r].float() - ref).abs().max().item()
            assert err <= 0.08 * ref.abs().max().item(), err
    finally:
        world.destroy()


def test_ht_capacity_error_is_loud():
    from b200ep import B200EPError, LocalWorld

    world = LocalWorld(1, sym_bytes=16 << 20)
    try:
        x, ids, w = make_rank_inputs(4096, 7168, 256, 8, 1)
        with pytest.raises(B200EPError, match="capacity"):
            world.bufs[0].dispatch(x.to(DEV), ids.to(DEV), w.to(DEV), 256, quantize=True)
    finally:
        world.destroy()


def test_ht_one_call_dispatch_single_rank():
    """Buffer.dispatch without a precomputed layout goes through b200ep_ht_dispatch_all (one host call)."""
    from b200ep import Buffer

    T, H, E, K = 300, 1024, 32, 8
    buf = Buffer(None, sym_bytes=128 << 20, rank=0, world=1)
    try:
        for it in range(3):
            x, ids, w = make_rank_inputs(T + it, H, E, K, seed=90 + it, mask_frac=0.15)
            q, s = fp8.group_quant_fp8(bf16_bits(x))
            ref = dispatch.ht_dispatch([q], [ids.numpy()], [w.numpy()], E, scales=[s])
            rx, rs, ridx, rw, per_expert, h = buf.dispatch(x.to(DEV), ids.to(DEV), w.to(DEV), E, quantize=True)
            rsrc = buf.ht_views(h)[4]
            _check_dispatch([(rx, rs, ridx, rw, rsrc, per_expert, h)], ref, "quant", K)
            y = bits.f32_to_bf16(fp8.dequant_fp8(ref[0]["recv_x"], ref[0]["recv_x_scales"]))
            out = buf.combine(bits_to_bf16(y, DEV), h)
            want = combine.ht_combine([y], ref, [T + it])[0]
            assert np.array_equal(bf16_bits(out), want)
    finally:
        buf.destroy()


def test_ht_host_sync_free_dispatch_single_rank():
    """num_worst_tokens > 0: no host wait; counts stay on the device; first num_recv rows equal the
    synchronous result and combine gives the same output."""
    from b200ep import Buffer

    T, H, E, K = 200, 1024, 32, 8
    buf = Buffer(None, sym_bytes=128 << 20, rank=0, world=1)
    try:
        x, ids, w = make_rank_inputs(T, H, E, K, seed=123, mask_frac=0.2)
        ids[3] = -1  # a token with no expert at all: fewer rows than tokens
        q, s = fp8.group_quant_fp8(bf16_bits(x))
        ref = dispatch.ht_dispatch([q], [ids.numpy()], [w.numpy()], E, scales=[s])[0]
        rx, rs, ridx, rw, per_expert, h = buf.dispatch(x.to(DEV), ids.to(DEV), w.to(DEV), E, quantize=True,
                                                       num_worst_tokens=T)
        assert per_expert == [] and rx.shape == (T, H)
        torch.cuda.synchronize()
        M = int(h.num_recv_dev)
        assert M == ref["num_recv_tokens"] and M < T
        assert np.array_equal(h.recv_per_expert_dev.cpu().numpy(), ref["num_recv_tokens_per_expert"])
        assert np.array_equal(u8(rx)[:M], ref["recv_x"]) and np.array_equal(ridx.cpu().numpy()[:M], ref["recv_topk_idx"])
        y = torch.zeros(T, H, dtype=torch.bfloat16, device=DEV)
        yv = bits.f32_to_bf16(fp8.dequant_fp8(ref["recv_x"], ref["recv_x_scales"]))
        y[:M] = bits_to_bf16(yv, DEV)
        out = buf.combine(y, h)
        want = combine.ht_combine([yv], [ref], [T])[0]
        assert np.array_equal(bf16_bits(out), want)
    finally:
        buf.destroy()


@pytest.mark.parametrize("ue8m0", [False, True])
def test_ht_full_size_bit_exact_single_rank(ue8m0):
    """The instantiation bench.py times at EP=1 -- ht_dispatch_send_kernel<quantise, hidden 7168> through
    Buffer.dispatch (one C call) -- at the full BASELINE shape (4096 tokens, 256 experts, top-8), byte for
    byte against oracle/fp8.py + oracle/dispatch.py, then combine against oracle/combine.py; and the received
    bytes/scales against vLLM's own CUDA quantiser on the same rows (per_token_group_quant_fp8)."""
    import parity
    from b200ep import Buffer

    T, H, E, K = 4096, 7168, 256, 8
    buf = Buffer(None, sym_bytes=1 << 30, rank=0, world=1)
    try:
        for it, kind in enumerate(("uniform", "zipf")):
            parity.check_ht(buf, None, 1, 0, torch.device(DEV), [T - 3 * it], H, E, K, seed=900 + it, dist=kind, ue8m0=ue8m0,
                            mask_frac=0.05 * it)
        parity.check_ht(buf, None, 1, 0, torch.device(DEV), [T], H, E, K, seed=910, nosync=True, ue8m0=ue8m0)
        try:
            from vllm.model_executor.layers.quantization.utils.fp8_utils import per_token_group_quant_fp8
        except Exception:  # pragma: no cover
            per_token_group_quant_fp8 = None
        if per_token_group_quant_fp8 is not None:
            x, ids, w = make_rank_inputs(T, H, E, K, seed=920)
            xd = x.to(DEV)
            rx, rs, ridx, rw, per_expert, h = buf.dispatch(xd, ids.to(DEV), w.to(DEV), E, quantize=True, ue8m0=ue8m0)
            vq, vs = per_token_group_quant_fp8(xd, 128, use_ue8m0=ue8m0)
            torch.cuda.synchronize()
            assert h.num_recv == T  # one rank, no masked ids: every token, ascending
            assert torch.equal(rx.view(torch.uint8), vq.view(torch.uint8)), "fp8 bytes differ from vLLM's CUDA op"
            assert torch.equal(rs.view(torch.int32), vs.contiguous().view(torch.int32)), "scales differ from vLLM's CUDA op"
    finally:
        buf.destroy()
