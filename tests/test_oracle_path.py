"""Oracle dispatch/combine vs golden vectors produced by the installed vLLM
(tests/golden/*.npz, made by oracle/gen_golden.py) and vs its own invariants."""
import os

import numpy as np
import pytest
import torch

from oracle import bits, combine, dispatch, fp8, layout
from oracle.cpu_reference import cpu_dispatch_combine

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _bf16_t(u16):
    return torch.from_numpy(u16.view(np.int16).copy()).view(torch.bfloat16)


def test_port_matches_vllm_cpu_fused_moe_bit_exact():
    g = np.load(os.path.join(GOLD, "cpu_moe_granite.npz"))
    x, coef = _bf16_t(g["x"]), _bf16_t(g["coef"])
    ids, w = torch.from_numpy(g["topk_ids"]), torch.from_numpy(g["topk_weights"])
    for stable in (False, True):
        out = cpu_dispatch_combine(x, w, ids, 32, expert_fn=lambda e, r: r * coef[e], stable=stable)
        assert np.array_equal(out.view(torch.int16).numpy().view(np.uint16), g["out"])


@pytest.mark.parametrize("W", [1, 2, 4, 8])
def test_ht_pipeline_vs_vllm_cpu_golden(W):
    """dispatch -> per-expert scaling -> on-rank weighted partial sum (what vLLM's
    TopKWeightAndReduceContiguous does before the HT combine) -> combine, compared
    with the real cpu_fused_moe_torch output: max abs err <= 1e-2 * max|out|."""
    g = np.load(os.path.join(GOLD, "cpu_moe_granite.npz"))
    E, K = 32, 8
    x, ids, w = g["x"], g["topk_ids"], g["topk_weights"]
    T = x.shape[0]
    cuts = np.linspace(0, T, W + 1).astype(int)
    xs = [x[cuts[i]:cuts[i + 1]] for i in range(W)]
    idl = [ids[cuts[i]:cuts[i + 1]] for i in range(W)]
    wl = [w[cuts[i]:cuts[i + 1]] for i in range(W)]
    disp = dispatch.ht_dispatch(xs, idl, wl, E)
    coef = bits.bf16_to_f32(g["coef"])
    ys = []
    for d in range(W):
        rx = bits.bf16_to_f32(disp[d]["recv_x"])
        lid, lw = disp[d]["recv_topk_idx"], disp[d]["recv_topk_weights"]
        acc = np.zeros_like(rx)
        for k in range(K):
            m = lid[:, k] >= 0
            ge = lid[m, k] + d * (E // W)
            eo = bits.bf16_to_f32(bits.f32_to_bf16(rx[m] * coef[ge][:, None]))  # expert output is bf16
            acc[m] += eo * lw[m, k][:, None]
        ys.append(bits.f32_to_bf16(acc))
    outs = combine.ht_combine(ys, disp, [c.shape[0] for c in xs])
    got = bits.bf16_to_f32(np.concatenate(outs, axis=0))
    ref = bits.bf16_to_f32(g["out"])
    assert np.abs(got - ref).max() <= 1e-2 * np.abs(ref).max()
    # received rows are verbatim copies, in canonical order
    for d in range(W):
        pre = disp[d]["rank_prefix"]
        for s in range(W):
            t = disp[d]["recv_src_idx"][pre[s]:pre[s + 1]]
            assert (np.diff(t) > 0).all()
            assert np.array_equal(disp[d]["recv_x"][pre[s]:pre[s + 1]], xs[s][t])


@pytest.mark.parametrize("W", [1, 2, 4])
def test_ll_layout_vs_vllm_batched_prepare(W):
    g = np.load(os.path.join(GOLD, "batched_prepare.npz"))
    E, max_tok = int(g["E"]), int(g["max_tok"])
    x, ids = g["x"], g["topk_ids"]
    T = x.shape[0]
    cuts = np.linspace(0, T, W + 1).astype(int)
    xs = [x[cuts[i]:cuts[i + 1]] for i in range(W)]
    idl = [ids[cuts[i]:cuts[i + 1]] for i in range(W)]
    disp = dispatch.ll_dispatch(xs, idl, E, max_tok)
    for r in range(W):
        cnt = g[f"w{W}_r{r}_cnt"][: E // W]  # vLLM quirk: length-E vector, first E_local filled
        assert np.array_equal(disp[r]["recv_count"], cnt)
        ref = g[f"w{W}_r{r}_x"]
        for el in range(E // W):
            n = cnt[el]
            assert np.array_equal(disp[r]["packed_recv_x"][el, :n], ref[el, :n])


@pytest.mark.parametrize("W", [1, 2, 4, 8])
@pytest.mark.parametrize("dist", ["uniform", "one_rank", "masked"])
def test_layout_invariants(W, dist):
    rng = np.random.default_rng(W)
    E, K, T = 64, 8, 257
    if dist == "one_rank":
        ids = np.stack([rng.permutation(E // W)[:K] for _ in range(T)]) + (W - 1) * (E // W)
    else:
        ids = np.stack([rng.permutation(E)[:K] for _ in range(T)])
    if dist == "masked":
        ids[rng.random(ids.shape) < 0.3] = -1
        ids[5] = -1
    tpr, tpe, inr, off = layout.dispatch_layout(ids, E, W)
    assert tpe.sum() == (ids >= 0).sum()
    assert np.array_equal(tpr, inr.sum(0))
    for d in range(W):
        rows = np.nonzero(inr[:, d])[0]
        assert np.array_equal(off[rows, d], np.arange(rows.size))
        assert (off[inr[:, d] == 0, d] == -1).all()
    # same counts as the head of cpu_fused_moe_torch (scatter_ + sum), on unmasked ids
    if dist != "masked":
        t = torch.from_numpy(ids)
        cn = t.new_zeros((T, E)); cn.scatter_(1, t, 1)
        assert np.array_equal(cn.sum(0).numpy(), tpe)


def test_empty_and_ragged():
    E, K, H = 16, 2, 128
    xs = [np.zeros((0, H), np.uint16), np.arange(3 * H, dtype=np.uint16).reshape(3, H)]
    ids = [np.zeros((0, K), np.int64), np.array([[0, 9], [8, 15], [1, 2]])]
    ws = [np.zeros((0, K), np.float32), np.ones((3, K), np.float32)]
    disp = dispatch.ht_dispatch(xs, ids, ws, E)
    assert disp[0]["num_recv_tokens"] == 2 and disp[1]["num_recv_tokens"] == 2
    assert np.array_equal(disp[0]["recv_src_idx"], [0, 2])
    assert np.array_equal(disp[1]["recv_topk_idx"], [[-1, 1], [0, 7]])
    outs = combine.ht_combine([d["recv_x"] for d in disp], disp, [0, 3])
    assert outs[0].shape == (0, H)
    # token 0 went to both ranks -> 2x, token 1 only to rank 1, token 2 only to rank 0
    got = bits.bf16_to_f32(outs[1]); src = bits.bf16_to_f32(xs[1])
    assert np.array_equal(got[1], src[1]) and np.array_equal(got[2], src[2])
    assert np.array_equal(got[0], bits.bf16_to_f32(bits.f32_to_bf16(2 * src[0])))


def test_ll_roundtrip_identity():
    rng = np.random.default_rng(0)
    W, E, K, H, T, mt = 4, 32, 4, 128, 9, 16
    xs = [bits.f32_to_bf16(rng.standard_normal((T - s, H)).astype(np.float32)) for s in range(W)]
    ids = [np.stack([rng.permutation(E)[:K] for _ in range(T - s)]) for s in range(W)]
    ws = [rng.random((T - s, K)).astype(np.float32) for s in range(W)]
    disp = dispatch.ll_dispatch(xs, ids, E, mt)
    ys = [d["packed_recv_x"] for d in disp]
    outs = combine.ll_combine(ys, ids, ws, disp, E)
    for s in range(W):
        ref = bits.bf16_to_f32(xs[s]) * ws[s].sum(1, keepdims=True)
        assert np.abs(bits.bf16_to_f32(outs[s]) - ref).max() <= 1e-2 * max(1.0, np.abs(ref).max())
    q, sc = fp8.group_quant_fp8(xs[0])
    assert np.abs(fp8.dequant_fp8(q, sc) - bits.bf16_to_f32(xs[0])).max() < 0.3


def test_fp8_oracle_matches_vllm_cuda_op_golden():
    """tests/golden/vllm_fp8_quant_gpu.npz was produced on a B200 by vLLM's CUDA quantiser
    (oracle/gen_golden_gpu.py): the oracle's "div" arithmetic must reproduce its bytes and scales,
    with and without UE8M0 -- this is what pins SURVEY H1 for the CPU suite."""
    g = np.load(os.path.join(GOLD, "vllm_fp8_quant_gpu.npz"))
    for ue in (0, 1):
        q, s = fp8.group_quant_fp8(g["x"], formula="div", ue8m0=bool(ue))
        assert np.array_equal(s.view(np.uint32), g[f"s_ue{ue}"].view(np.uint32)), f"scales (ue8m0={ue})"
        assert np.array_equal(q, g[f"q_ue{ue}"]), f"fp8 bytes (ue8m0={ue})"
    # and the other two named arithmetics do NOT reproduce it (they are not interchangeable)
    qr, sr = fp8.group_quant_fp8(g["x"], formula="recip")
    assert (sr.view(np.uint32) != g["s_ue0"].view(np.uint32)).any()
