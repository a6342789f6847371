"""Per-token-group FP8 (e4m3fn) quantisation, CPU restatement (test infrastructure).

Follows vLLM ``model_executor/layers/quantization/utils/fp8_utils.py``:
``per_token_group_quant_fp8`` (:510-591), the Triton restatement
``_per_token_group_quant_fp8`` (:94-149), eps = 1e-10 (:513) and the e4m3 range
(+-448, ``quant_utils.py:27-35``).  vLLM reaches it from the dispatch path at
``prepare_finalize/deepep_ht.py:282-289`` (before the HT dispatch) and
``deepep_ll.py:211-217`` (after the LL dispatch).

Three arithmetics exist for the same nominal formula and they do NOT produce the
same bytes, so each is named:

``"div"``     s = max(amax, eps) / 448      ; q = sat(x / s)     -- vLLM's CUDA op
              (IEEE division in its SASS); THE parity target of the CUDA kernels.
``"recip"``   s = max(amax, eps) * (1/448)  ; q = sat(x / s)     -- vLLM's Triton
              fallback (fp8_utils.py:143-147).
``"mulinv"``  r = 448 / max(amax, eps); q = sat(x * r); s = 1 / r -- the shape of
              DeepEP's in-kernel low-latency quantiser (recalled, not pinned).

``ue8m0=True`` rounds the scale up to a power of two, ``s = 2**ceil(log2 s)``
(fp8_utils.py:145), computed here exactly on the float32 bit pattern.
"""

from __future__ import annotations

import numpy as np

from .bits import E4M3_MAX, bf16_to_f32, f32_to_e4m3, f32_to_e4m3_fast

EPS = np.float32(1e-10)
GROUP = 128


def ceil_pow2_f32(s: np.ndarray) -> np.ndarray:
    """Smallest power of two >= max(s, 1e-10) for positive normal float32 s, bit-exact.
    (vLLM's CUDA op computes exp2f(ceilf(log2f(fmaxf(fabsf(s), 1e-10f)))): the scale is clamped a
    second time in the UE8M0 branch.)"""
    s = np.maximum(np.ascontiguousarray(s, dtype=np.float32), EPS)
    u = s.view(np.uint32)
    mant = u & np.uint32(0x007FFFFF)
    exp = u & np.uint32(0x7F800000)
    up = np.where(mant != 0, exp + np.uint32(0x00800000), exp)
    return up.astype(np.uint32).view(np.float32)


def group_quant_fp8(
    x: np.ndarray,
    *,
    x_is_bf16_bits: bool = True,
    group: int = GROUP,
    formula: str = "div",
    ue8m0: bool = False,
    fast: bool = False,
) -> tuple[np.ndarray, np.ndarray]:
    """Quantise rows of ``x`` in groups of ``group`` elements.

    x: [T, H] bf16 bit patterns (uint16) or float32 values.
    Returns (q uint8 [T, H] e4m3fn bit patterns, s float32 [T, H // group]).
    ``fast`` selects torch's CPU cast for the final fp32 -> e4m3 step (see
    ``bits.f32_to_e4m3_fast``); everything else is the same numpy arithmetic.
    """
    xf = bf16_to_f32(x) if x_is_bf16_bits else np.ascontiguousarray(x, np.float32)
    T, H = xf.shape
    assert H % group == 0
    g = xf.reshape(T, H // group, group)
    amax = np.maximum(np.max(np.abs(g), axis=2), EPS).astype(np.float32)
    f448 = np.float32(E4M3_MAX)
    if formula == "div":
        s = (amax / f448).astype(np.float32)
        if ue8m0:
            s = ceil_pow2_f32(s)
        y = (g / s[:, :, None]).astype(np.float32)
    elif formula == "recip":
        s = (amax * (np.float32(1.0) / f448)).astype(np.float32)
        if ue8m0:
            s = ceil_pow2_f32(s)
        y = (g / s[:, :, None]).astype(np.float32)
    elif formula == "mulinv":
        r = (f448 / amax).astype(np.float32)
        s = (np.float32(1.0) / r).astype(np.float32)
        if ue8m0:
            s = ceil_pow2_f32(s)
            r = (np.float32(1.0) / s).astype(np.float32)
        y = (g * r[:, :, None]).astype(np.float32)
    else:
        raise ValueError(formula)
    q = (f32_to_e4m3_fast if fast else f32_to_e4m3)(y.reshape(T, H))
    return q, s


def dequant_fp8(q: np.ndarray, s: np.ndarray, group: int = GROUP) -> np.ndarray:
    """fp32 dequantisation, vLLM ``deepep_ll.py:34-49`` (``dequant_fp8``)."""
    from .bits import e4m3_to_f32

    T, H = q.shape
    v = e4m3_to_f32(q).reshape(T, H // group, group)
    return (v * s[:, :, None]).astype(np.float32).reshape(T, H)


def pack_ue8m0(s: np.ndarray) -> np.ndarray:
    """Power-of-two float32 scales [T, G] -> packed int32 [T, ceil(G/4)], one
    biased exponent byte per scale, little-endian within the word (vLLM
    ``fp8_utils.py:643-714`` packed-UE8M0 layout)."""
    s = np.ascontiguousarray(s, dtype=np.float32)
    T, G = s.shape
    e = ((s.view(np.uint32) >> np.uint32(23)) & np.uint32(0xFF)).astype(np.uint8)
    Gp = (G + 3) // 4 * 4
    pad = np.zeros((T, Gp), np.uint8)
    pad[:, :G] = e
    return pad.view(np.int32).reshape(T, Gp // 4)
