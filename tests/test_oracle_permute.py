"""CPU checks of the expert-grouped permute oracle (oracle/permute.py): the fused multiply-add restatement against exact
rational arithmetic, and the layout contract of vLLM's ep_scatter / ep_gather (fused_moe/deep_gemm_utils.py:56-344)."""
from fractions import Fraction
import math

import numpy as np

from oracle import bits, permute


def _exact_fma(a, b, c):
    f = Fraction(float(a)) * Fraction(float(b)) + Fraction(float(c))
    if f == 0:
        return np.float32(0)
    sign = -1 if f < 0 else 1
    f = abs(f)
    e = math.floor(math.log2(f))
    while Fraction(2) ** e > f:
        e -= 1
    while Fraction(2) ** (e + 1) <= f:
        e += 1
    q = Fraction(2) ** (e - 23)
    n = f / q
    fl = n.numerator // n.denominator
    rem = n - fl
    if rem > Fraction(1, 2) or (rem == Fraction(1, 2) and fl % 2 == 1):
        fl += 1
    return np.float32(sign * float(fl * q))


def test_fma32_is_a_single_rounding():
    rng = np.random.default_rng(0)
    a = rng.standard_normal(1500).astype(np.float32)
    b = rng.random(1500).astype(np.float32)
    c = rng.standard_normal(1500).astype(np.float32)
    # double-rounding traps: the float64 sum lands exactly on a float32 tie, the exact sum just beside it
    eps = np.float32(2.0**-23)
    ta = np.full(64, 1 + eps, np.float32)
    tb = (np.float32(2.0**-24) * (1 - eps)) * np.ones(64, np.float32)   # a*b = 2^-24 - 2^-70
    tc = (1 + (2 * np.arange(64) + 1) * 2.0**-23).astype(np.float32)    # odd significands: ties would round up
    A = np.concatenate([a, ta, -ta, ta])
    B = np.concatenate([b, tb, tb, -tb])
    Cc = np.concatenate([c, tc, -tc, tc])
    got = permute.fma32(A, B, Cc)
    naive = (A.astype(np.float64) * B.astype(np.float64) + Cc.astype(np.float64)).astype(np.float32)
    want = np.array([_exact_fma(x, y, z) for x, y, z in zip(A, B, Cc)], np.float32)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert (naive.view(np.uint32) != want.view(np.uint32)).any(), "the trap inputs no longer exercise double rounding"


def _case(M, K, E, H, seed, mask=0.2):
    rng = np.random.default_rng(seed)
    ids = np.stack([rng.permutation(E + 3)[:K] for _ in range(M)]).astype(np.int64) if M else np.zeros((0, K), np.int64)
    ids[ids >= E] = -1                      # experts that live on other ranks
    ids[rng.random((M, K)) < mask] = -1
    x = rng.integers(0, 256, (M, H), dtype=np.uint8)
    s = rng.random((M, H // 128)).astype(np.float32)
    w = rng.random((M, K)).astype(np.float32)
    return x, s, ids, w


def test_permute_layout_contract():
    for M, K, E, seed in [(0, 8, 4, 0), (1, 8, 32, 1), (300, 8, 32, 2), (1000, 6, 5, 3), (257, 1, 1, 4)]:
        x, s, ids, w = _case(M, K, E, 256, seed)
        p = permute.moe_permute(x, s, ids, E, alignment=128)
        start, total = permute.expert_starts(p["per_expert"], 128)
        assert total == p["m_sum"] and total % 128 == 0 and np.all(start % 128 == 0)
        # every entry with a local expert got exactly one row; rows of an expert are contiguous, in received order
        for e in range(E):
            rows = np.nonzero(p["expert_ids"] == e)[0]
            assert rows.size == p["per_expert"][e]
            if rows.size:
                assert rows[0] == start[e] and rows[-1] == start[e] + rows.size - 1
                src = np.nonzero((ids == e).any(1))[0]
                assert np.array_equal(p["out_x"][rows], x[src]) and np.array_equal(p["out_scales"][rows], s[src])
        ok = ids >= 0
        assert np.all(p["inv_perm"][~ok] == -1)
        assert np.unique(p["inv_perm"][ok]).size == int(ok.sum())
        assert np.all(p["expert_ids"][p["inv_perm"][ok]] == ids[ok])
        assert np.all(p["expert_ids"][~p["valid"]] == -1)


def test_unpermute_reduce_matches_a_direct_sum():
    M, K, E, H = 200, 8, 16, 256
    x, s, ids, w = _case(M, K, E, H, 7)
    p = permute.moe_permute(x, s, ids, E)
    rng = np.random.default_rng(1)
    y = bits.f32_to_bf16(rng.standard_normal((p["m_sum"], H)).astype(np.float32))
    z = permute.moe_unpermute_reduce(y, ids, w, p["inv_perm"], E)
    yf = bits.bf16_to_f32(y).astype(np.float64)
    ref = np.zeros((M, H))
    for m in range(M):
        for k in range(K):
            if ids[m, k] >= 0:
                ref[m] += w[m, k] * yf[p["inv_perm"][m, k]]
    got = bits.bf16_to_f32(z).astype(np.float64)
    assert np.all(np.abs(got - ref) <= 2.0**-8 * np.maximum(np.abs(ref), 1e-3) + 1e-6)
    # rows with no local expert come back as zeros
    none = ~(ids >= 0).any(1)
    assert none.any() or True
    assert np.all(z[none] == 0)


def test_aligned_rows_matches_vllm_compute_aligned_m():
    """Host-side sizing of the grouped layout == vLLM's compute_aligned_M (fused_moe/deep_gemm_utils.py:26-44) in both of
    its modes: exact from the per-expert counts, worst case without them; and == the oracle's total."""
    import pytest

    dgu = pytest.importorskip("vllm.model_executor.layers.fused_moe.deep_gemm_utils")
    mk = pytest.importorskip("vllm.model_executor.layers.fused_moe.modular_kernel")
    import torch

    from b200ep import Buffer

    rng = np.random.default_rng(3)
    for E, M, K in ((32, 21714, 8), (256, 4096, 8), (36, 100, 6), (1, 0, 1)):
        cnt = rng.integers(0, 900, E).astype(np.int32)
        meta = mk.ExpertTokensMetadata(expert_num_tokens=torch.from_numpy(cnt), expert_num_tokens_cpu=torch.from_numpy(cnt))
        want = dgu.compute_aligned_M(M, K, E, 128, meta)
        assert Buffer.aligned_rows(M, K, E, 128, cnt.tolist()) == want == permute.expert_starts(cnt, 128)[1]
        assert Buffer.aligned_rows(M, K, E, 128) == dgu.compute_aligned_M(M, K, E, 128, None)
