"""GPU parity: low-latency dispatch + combine vs the CPU oracle, through the C ABI."""
import numpy as np
import pytest
import torch

from _util import bf16_bits, bits_to_bf16, make_rank_inputs, u8
from oracle import bits, combine, dispatch, fp8

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("W", [1, 2, 4, 8])
@pytest.mark.parametrize("use_fp8", [True, False])
def test_ll_dispatch_combine_bit_exact(W, use_fp8):
    from b200ep import LocalWorld

    H, E, K, max_tok = 2048, 64, 8, 48
    world = LocalWorld(W, sym_bytes=256 << 20)
    try:
        for it, (dist, mask) in enumerate([("uniform", 0.0), ("one_rank", 0.0), ("zipf", 0.25)]):
            Ts = [max_tok - 5 * r if r != 1 else 0 for r in range(W)]
            if W == 1:
                Ts = [max_tok]
            xs, ids, ws = [], [], []
            for r in range(W):
                x, i, w = make_rank_inputs(Ts[r], H, E, K, 1000 * it + r + 17 * W, dist=dist, world=W, mask_frac=mask)
                xs.append(x); ids.append(i); ws.append(w)
            xb = [bf16_bits(x) for x in xs]
            if use_fp8:
                qs = [fp8.group_quant_fp8(b, formula="div") for b in xb]
                ref = dispatch.ll_dispatch([q for q, _ in qs], [i.numpy() for i in ids], E, max_tok,
                                           scales=[s for _, s in qs])
            else:
                ref = dispatch.ll_dispatch(xb, [i.numpy() for i in ids], E, max_tok)
            res = world.low_latency_dispatch([x.to(DEV) for x in xs], [i.to(DEV) for i in ids], max_tok, E,
                                             use_fp8=use_fp8)
            ys_np = []
            for d, (px, ps, cnt, h) in enumerate(res):
                o = ref[d]
                assert np.array_equal(cnt.cpu().numpy(), o["recv_count"])
                assert np.array_equal(h.layout_range.cpu().numpy(), o["layout_range"])
                pxh = u8(px) if use_fp8 else bf16_bits(px)
                psh = ps.cpu().numpy() if use_fp8 else None
                sih = h.src_info.cpu().numpy()
                y = np.zeros((E // W, W * max_tok, H), np.float32)
                for el in range(E // W):
                    n = int(o["recv_count"][el])
                    assert np.array_equal(pxh[el, :n], o["packed_recv_x"][el, :n])
                    assert np.array_equal(sih[el, :n], o["src_info"][el, :n])
                    if use_fp8:
                        assert np.array_equal(psh[el, :n].view(np.uint32), o["packed_recv_scales"][el, :n].view(np.uint32))
                        y[el, :n] = fp8.dequant_fp8(o["packed_recv_x"][el, :n], o["packed_recv_scales"][el, :n])
                    else:
                        y[el, :n] = bits.bf16_to_f32(o["packed_recv_x"][el, :n])
                    y[el, :n] *= 1.0 + 0.5 * el
                ys_np.append(bits.f32_to_bf16(y.reshape(-1, H)).reshape(E // W, W * max_tok, H))
            outs = world.low_latency_combine([bits_to_bf16(y, DEV) for y in ys_np], [i.to(DEV) for i in ids],
                                             [w.to(DEV) for w in ws], [r[-1] for r in res])
            ref_out = combine.ll_combine(ys_np, [i.numpy() for i in ids], [w.numpy() for w in ws], ref, E)
            for r in range(W):
                assert np.array_equal(bf16_bits(outs[r]), ref_out[r]), f"ll combine differs on rank {r}"
    finally:
        world.destroy()


def test_ll_decode_shape_roundtrip():
    """BASELINE config #3 shape at W=1: 128 tokens, hidden 7168, 256 experts, top-8, fp8."""
    from b200ep import LocalWorld

    T, H, E, K = 128, 7168, 256, 8
    world = LocalWorld(1, sym_bytes=2 << 30)
    try:
        x, ids, w = make_rank_inputs(T, H, E, K, 3)
        (px, ps, cnt, h), = world.low_latency_dispatch([x.to(DEV)], [ids.to(DEV)], 128, E, use_fp8=True)
        assert int(cnt.sum()) == T * K
        y = (px.float().view(E, 128, H // 128, 128) * ps[..., None]).view(E, 128, H).to(torch.bfloat16)
        out, = world.low_latency_combine([y], [ids.to(DEV)], [w.to(DEV)], [h])
        ref = x.to(DEV).float()  # weights sum to 1
        assert (out.float() - ref).abs().max().item() <= 0.08 * ref.abs().max().item()
    finally:
        world.destroy()


@pytest.mark.parametrize("hook", [False, True])
def test_ll_single_launch_paths_single_rank(hook):
    """The production entry points (count+send in one launch, optional recv hook, one-launch combine)
    on one rank, against the oracle -- the phase-by-phase kernels are covered above."""
    from b200ep import Buffer

    T, H, E, K, max_tok = 57, 2048, 32, 8, 64
    buf = Buffer(None, sym_bytes=256 << 20, rank=0, world=1)
    try:
        for it in range(3):  # repeated calls: epochs and both slots
            x, ids, w = make_rank_inputs(T - it, H, E, K, seed=40 + it, mask_frac=0.1 * it)
            q, s = fp8.group_quant_fp8(bf16_bits(x))
            ref = dispatch.ll_dispatch([q], [ids.numpy()], E, max_tok, scales=[s])[0]
            px, ps, cnt, h, hk = buf.low_latency_dispatch(x.to(DEV), ids.to(DEV), max_tok, E, use_fp8=True,
                                                          return_recv_hook=hook)
            if hook:
                hk()
            torch.cuda.synchronize()
            assert np.array_equal(cnt.cpu().numpy(), ref["recv_count"])
            y = np.zeros((E, max_tok, H), np.float32)
            for el in range(E):
                n = int(ref["recv_count"][el])
                assert np.array_equal(u8(px)[el, :n], ref["packed_recv_x"][el, :n])
                assert np.array_equal(ps.cpu().numpy()[el, :n], ref["packed_recv_scales"][el, :n])
                assert np.array_equal(h.src_info.cpu().numpy()[el, :n], ref["src_info"][el, :n])
                y[el, :n] = fp8.dequant_fp8(ref["packed_recv_x"][el, :n], ref["packed_recv_scales"][el, :n]) * (1 + el)
            yb = bits.f32_to_bf16(y.reshape(-1, H)).reshape(E, max_tok, H)
            out, chk = buf.low_latency_combine(bits_to_bf16(yb, DEV), ids.to(DEV), w.to(DEV), h, return_recv_hook=hook)
            if hook:
                chk()
            torch.cuda.synchronize()
            want = combine.ll_combine([yb], [ids.numpy()], [w.numpy()], [ref], E)[0]
            assert np.array_equal(bf16_bits(out), want)
    finally:
        buf.destroy()


def test_ll_eplb_expert_count_world8():
    """guides/wide-ep-lws decode.yaml:114-118 runs with 32 redundant experts: 288 physical experts,
    36 per rank at EP=8.  Ids are physical ids here (vLLM maps them before the call, deepep_ll.py:291)."""
    from b200ep import LocalWorld

    W, H, E, K, max_tok = 8, 256, 288, 8, 16
    world = LocalWorld(W, sym_bytes=64 << 20)
    try:
        ins = [make_rank_inputs(max_tok - r, H, E, K, 500 + r) for r in range(W)]
        xs, ids = [i[0] for i in ins], [i[1] for i in ins]
        qs = [fp8.group_quant_fp8(bf16_bits(x)) for x in xs]
        ref = dispatch.ll_dispatch([q for q, _ in qs], [i.numpy() for i in ids], E, max_tok, scales=[s for _, s in qs])
        res = world.low_latency_dispatch([x.to(DEV) for x in xs], [i.to(DEV) for i in ids], max_tok, E, use_fp8=True)
        for d, (px, ps, cnt, h) in enumerate(res):
            assert px.shape[0] == 36
            assert np.array_equal(cnt.cpu().numpy(), ref[d]["recv_count"])
            pxh = u8(px)
            for el in range(36):
                n = int(ref[d]["recv_count"][el])
                assert np.array_equal(pxh[el, :n], ref[d]["packed_recv_x"][el, :n])
    finally:
        world.destroy()


@pytest.mark.parametrize("W,H", [(1, 7168), (2, 2048), (4, 640)])
def test_ll_packed_ue8m0_scales_written_by_the_kernel(W, H):
    """use_ue8m0 (DeepEP) / packed UE8M0 (vLLM fp8_utils.py:643-714): power-of-two scales travel as exponent
    bytes, four per int32, slot index fastest.  Bytes must equal the oracle's pack of the oracle's UE8M0 scales,
    the payload must equal the UE8M0 quantisation, and the view must have the MN-major strides DeepGEMM checks."""
    from b200ep import LocalWorld

    E, K, max_tok = 32, 8, 40
    world = LocalWorld(W, sym_bytes=256 << 20)
    try:
        ins = [make_rank_inputs(max_tok - 3 * r, H, E, K, 77 + r, mask_frac=0.1) for r in range(W)]
        xs, ids = [i[0] for i in ins], [i[1] for i in ins]
        qs = [fp8.group_quant_fp8(bf16_bits(x), ue8m0=True) for x in xs]
        ref = dispatch.ll_dispatch([q for q, _ in qs], [i.numpy() for i in ids], E, max_tok, scales=[s for _, s in qs])
        res = world.low_latency_dispatch([x.to(DEV) for x in xs], [i.to(DEV) for i in ids], max_tok, E, use_fp8=True, ue8m0=2)
        g4, cap = (H // 128 + 3) // 4, W * max_tok
        for d, (px, ps, cnt, h) in enumerate(res):
            assert ps.dtype == torch.int32 and tuple(ps.shape) == (E // W, cap, g4)
            assert ps.stride(1) == 1 and ps.stride(2) == (cap + 3) // 4 * 4 and ps.stride(0) == g4 * ps.stride(2)
            assert np.array_equal(cnt.cpu().numpy(), ref[d]["recv_count"])
            pxh, psh = u8(px), ps.cpu().numpy()
            for el in range(E // W):
                n = int(ref[d]["recv_count"][el])
                assert np.array_equal(pxh[el, :n], ref[d]["packed_recv_x"][el, :n])
                assert np.array_equal(psh[el, :n], fp8.pack_ue8m0(ref[d]["packed_recv_scales"][el, :n]))
    finally:
        world.destroy()


def test_ll_packed_ue8m0_matches_vllm_packed_quantiser():
    """Layout AND bytes against vLLM's own packed-UE8M0 quantiser for DeepGEMM
    (per_token_group_quant_fp8_packed_for_deepgemm, fp8_utils.py:643-714) on the same rows: with one rank and
    one expert per token every token lands in slot order, so the two [rows, H/512] int32 views must be equal."""
    try:
        from vllm.model_executor.layers.quantization.utils.fp8_utils import per_token_group_quant_fp8_packed_for_deepgemm
    except Exception as e:  # pragma: no cover
        pytest.skip(f"vLLM packed quantiser not importable: {e}")
    from b200ep import LocalWorld

    T, H, E = 96, 7168, 1
    world = LocalWorld(1, sym_bytes=256 << 20)
    try:
        x, _, _ = make_rank_inputs(T, H, 8, 1, 5)
        ids = torch.zeros(T, 1, dtype=torch.int64)
        (px, ps, cnt, h), = world.low_latency_dispatch([x.to(DEV)], [ids.to(DEV)], T, E, use_fp8=True, ue8m0=2)
        try:
            vq, vs = per_token_group_quant_fp8_packed_for_deepgemm(x.to(DEV), 128, use_ue8m0=True)
        except Exception as e:  # pragma: no cover  (op missing from this vLLM build)
            pytest.skip(f"vLLM packed quantiser not runnable here: {e}")
        torch.cuda.synchronize()
        assert int(cnt[0]) == T
        assert torch.equal(px[0, :T].view(torch.uint8), vq.view(torch.uint8))
        assert tuple(vs.shape) == tuple(ps[0, :T].shape) and vs.stride(0) == ps.stride(1) == 1
        assert torch.equal(ps[0, :T].contiguous(), vs.contiguous())
    finally:
        world.destroy()


def test_ll_wide_rows_and_eplb_combine():
    """(a) hidden 16384: a combine CTA then has fewer warps than rows per grab -- every grabbed row must still be
    pushed; (b) 288 physical experts (36 per rank at EP=8, guides/wide-ep-lws decode.yaml:114-118) through the
    combine as well."""
    from b200ep import LocalWorld

    for W, H, E, K, max_tok in ((2, 16384, 16, 4, 24), (8, 256, 288, 8, 16)):
        world = LocalWorld(W, sym_bytes=(512 << 20) if H > 4096 else (64 << 20))
        try:
            ins = [make_rank_inputs(max_tok - r, H, E, K, 900 + r + W) for r in range(W)]
            xs, ids, ws = [i[0] for i in ins], [i[1] for i in ins], [i[2] for i in ins]
            xb = [bf16_bits(x) for x in xs]
            ref = dispatch.ll_dispatch(xb, [i.numpy() for i in ids], E, max_tok)
            res = world.low_latency_dispatch([x.to(DEV) for x in xs], [i.to(DEV) for i in ids], max_tok, E, use_fp8=False)
            ys = []
            for d, (px, ps, cnt, h) in enumerate(res):
                assert np.array_equal(cnt.cpu().numpy(), ref[d]["recv_count"])
                y = bits.bf16_to_f32(ref[d]["packed_recv_x"]) * (1.0 + 0.5 * np.arange(E // W, dtype=np.float32))[:, None, None]
                ys.append(bits.f32_to_bf16(y.reshape(-1, H)).reshape(E // W, W * max_tok, H))
            outs = world.low_latency_combine([bits_to_bf16(y, DEV) for y in ys], [i.to(DEV) for i in ids],
                                             [w.to(DEV) for w in ws], [r[-1] for r in res])
            want = combine.ll_combine(ys, [i.numpy() for i in ids], [w.numpy() for w in ws], ref, E)
            for r in range(W):
                assert np.array_equal(bf16_bits(outs[r]), want[r]), f"W={W} H={H}: ll combine differs on rank {r}"
        finally:
            world.destroy()
