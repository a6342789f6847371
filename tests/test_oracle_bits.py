"""Oracle number formats vs torch's CPU casts and the committed golden vectors."""
import os

import numpy as np
import pytest
import torch

from oracle import bits, fp8

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_casts_golden():
    g = np.load(os.path.join(GOLD, "casts.npz"))
    v = g["f32"]
    assert np.array_equal(bits.f32_to_e4m3(v), g["e4m3_sat"])
    assert np.array_equal(bits.f32_to_bf16(v), g["bf16"])


def test_e4m3_all_codes_roundtrip():
    codes = np.arange(256, dtype=np.uint8)
    finite = (codes & 0x7F) != 0x7F
    f = bits.e4m3_to_f32(codes)
    ref = torch.from_numpy(codes).view(torch.float8_e4m3fn).to(torch.float32).numpy()
    assert np.array_equal(f[finite], ref[finite])
    back = bits.f32_to_e4m3(f[finite])
    assert np.array_equal(back, codes[finite])


@pytest.mark.parametrize("seed", [0, 1])
def test_e4m3_random_vs_torch(seed):
    rng = np.random.default_rng(seed)
    v = (rng.standard_normal(200000) * np.exp(rng.uniform(-12, 7, 200000))).astype(np.float32)
    ref = torch.from_numpy(v).clamp(-448.0, 448.0).to(torch.float8_e4m3fn).view(torch.uint8).numpy()
    assert np.array_equal(bits.f32_to_e4m3(v), ref)


def test_bf16_random_vs_torch():
    rng = np.random.default_rng(5)
    v = rng.standard_normal(100000).astype(np.float32)
    ref = torch.from_numpy(v).to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
    assert np.array_equal(bits.f32_to_bf16(v), ref)
    assert np.array_equal(bits.bf16_to_f32(ref), torch.from_numpy(v).to(torch.bfloat16).float().numpy())


@pytest.mark.parametrize("formula", ["div", "recip", "mulinv"])
def test_group_quant_vs_torch_restatement(formula):
    """Independent torch restatement of the three named arithmetics."""
    g = torch.Generator().manual_seed(11)
    x = (torch.randn(64, 512, generator=g) * (torch.rand(64, 1, generator=g) * 4)).to(torch.bfloat16)
    x[3] = 0  # all-zero groups exercise eps
    xb = x.view(torch.int16).numpy().view(np.uint16)
    q, s = fp8.group_quant_fp8(xb, formula=formula)
    xf = x.float().view(64, 4, 128)
    amax = xf.abs().amax(dim=2).clamp(min=1e-10)
    if formula == "div":
        ts = amax / 448.0
        y = xf / ts[..., None]
    elif formula == "recip":
        ts = amax * torch.tensor(1.0 / 448.0, dtype=torch.float32)
        y = xf / ts[..., None]
    else:
        # NB: torch evaluates `scalar / tensor` as reciprocal(tensor) * scalar; use tensor / tensor
        r = torch.full_like(amax, 448.0) / amax
        ts = torch.ones_like(r) / r
        y = xf * r[..., None]
    tq = y.clamp(-448, 448).to(torch.float8_e4m3fn).view(torch.uint8).view(64, 512).numpy()
    assert np.array_equal(s, ts.numpy())
    assert np.array_equal(q, tq)


def test_formulas_differ():
    """the three arithmetics are NOT interchangeable; keep it visible."""
    g = torch.Generator().manual_seed(0)
    x = torch.randn(256, 1024, generator=g).to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
    qd, sd = fp8.group_quant_fp8(x, formula="div")
    qr, sr = fp8.group_quant_fp8(x, formula="recip")
    assert (sd != sr).any()


def test_ue8m0_scale_is_pow2_and_covers():
    g = torch.Generator().manual_seed(2)
    x = torch.randn(32, 256, generator=g).to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
    _, s0 = fp8.group_quant_fp8(x, formula="div")
    q, s = fp8.group_quant_fp8(x, formula="div", ue8m0=True)
    assert ((s.view(np.uint32) & 0x7FFFFF) == 0).all()
    assert (s >= s0).all() and (s < 2 * s0).all()
    packed = fp8.pack_ue8m0(s)
    assert packed.shape == (32, 1)
    assert np.array_equal(packed.view(np.uint8).reshape(32, 4)[:, :2], (s.view(np.uint32) >> 23).astype(np.uint8))
    # exact powers of two stay put
    assert np.array_equal(fp8.ceil_pow2_f32(np.array([0.5, 1.0, 3.0], np.float32)), np.array([0.5, 1.0, 4.0], np.float32))
