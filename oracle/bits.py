"""Bit-level number formats used on the path, in plain numpy (test infrastructure).

bf16 is carried as uint16 and FP8 e4m3fn as uint8 so that every comparison in
the tests is an integer comparison.  The conversions are checked against
torch's own CPU casts (``tests/test_oracle_bits.py``) and against the golden
vectors made by ``oracle/gen_golden.py``.

Format constants follow vLLM ``model_executor/layers/quantization/utils/
quant_utils.py:27-35`` (e4m3fn: finite max 448, no infinities, 0x7f/0xff = NaN).
"""

from __future__ import annotations

import numpy as np

E4M3_MAX = 448.0


def bf16_to_f32(u16: np.ndarray) -> np.ndarray:
    """bf16 bit patterns (uint16) -> float32 values (exact)."""
    u16 = np.ascontiguousarray(u16, dtype=np.uint16)
    return (u16.astype(np.uint32) << np.uint32(16)).view(np.float32)


def f32_to_bf16(x: np.ndarray) -> np.ndarray:
    """float32 -> bf16 bit patterns, round-to-nearest-even (what ``Tensor.to(bfloat16)``
    and ``__float2bfloat16_rn`` do).  NaNs become the quiet NaN 0x7fc0 | sign."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    u = x.view(np.uint32)
    bias = np.uint32(0x7FFF) + ((u >> np.uint32(16)) & np.uint32(1))
    r = ((u + bias) >> np.uint32(16)).astype(np.uint16)
    nan = np.isnan(x)
    if nan.any():
        r = np.where(nan, ((u >> np.uint32(16)).astype(np.uint16) | np.uint16(0x0040)), r)
    return r


def f32_to_e4m3(x: np.ndarray) -> np.ndarray:
    """float32 -> e4m3fn bit patterns, round-to-nearest-even, saturate-to-finite.

    Same result as ``cvt.rn.satfinite.e4m3x2.f32`` (SASS ``F2FP.SATFINITE.E4M3``)
    and as ``clamp(x, -448, 448).to(torch.float8_e4m3fn)`` on the CPU.
    """
    x = np.ascontiguousarray(x, dtype=np.float32)
    sign = (x.view(np.uint32) >> np.uint32(24)).astype(np.uint8) & np.uint8(0x80)
    a = np.abs(x).astype(np.float64)
    nan = np.isnan(a)
    a = np.where(nan, 0.0, a)
    a = np.minimum(a, E4M3_MAX)  # satfinite: +-inf and anything above 448 -> 448
    # exponent of the value (floor(log2 a)), clamped to the subnormal exponent -6
    with np.errstate(divide="ignore"):
        _, e = np.frexp(a)  # a = m * 2**e, m in [0.5, 1)
    e = np.maximum(e.astype(np.int64) - 1, -6)
    quantum = np.ldexp(1.0, (e - 3).astype(np.int64))  # 3 mantissa bits
    q = np.rint(a / quantum) * quantum  # np.rint is round-half-even; exact in f64
    q = np.minimum(q, E4M3_MAX)
    # encode
    with np.errstate(divide="ignore"):
        _, e2 = np.frexp(q)
    e2 = e2.astype(np.int64) - 1
    normal = q >= 2.0**-6
    exp_field = np.where(normal, e2 + 7, 0)
    mant = np.where(
        normal,
        np.rint((q / np.ldexp(1.0, np.where(normal, e2, 0)) - 1.0) * 8.0),
        np.rint(q * 512.0),  # subnormal: m * 2**-9
    ).astype(np.int64)
    bits = ((exp_field << 3) | mant).astype(np.uint8)
    bits = np.where(nan, np.uint8(0x7F), bits)
    return bits | sign


def e4m3_to_f32(b: np.ndarray) -> np.ndarray:
    """e4m3fn bit patterns -> float32 (exact); 0x7f / 0xff -> NaN."""
    b = np.ascontiguousarray(b, dtype=np.uint8)
    if b.size > 4096:  # large inputs: table of the 256 codes, built with the arithmetic below
        return _e4m3_table()[b]
    s = np.where(b & 0x80, -1.0, 1.0)
    e = ((b >> 3) & 0x0F).astype(np.int64)
    m = (b & 0x07).astype(np.float64)
    v = np.where(e == 0, m * 2.0**-9, (1.0 + m / 8.0) * np.ldexp(1.0, e - 7))
    v = np.where((b & 0x7F) == 0x7F, np.nan, v)
    return (s * v).astype(np.float32)


_E4M3_TABLE = None


def _e4m3_table() -> np.ndarray:
    global _E4M3_TABLE
    if _E4M3_TABLE is None:
        _E4M3_TABLE = e4m3_to_f32(np.arange(256, dtype=np.uint8))
    return _E4M3_TABLE


def f32_to_e4m3_fast(x: np.ndarray) -> np.ndarray:
    """Same function as ``f32_to_e4m3`` on non-NaN inputs, through torch's CPU cast
    (``clamp(x, -448, 448).to(float8_e4m3fn)``), which is multi-threaded.  Used by the
    full-size parity checks where the pure-numpy version would take minutes;
    ``tests/test_oracle_bits.py`` holds the two equal on random and adversarial inputs."""
    import torch

    t = torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32))
    return t.clamp(-E4M3_MAX, E4M3_MAX).to(torch.float8_e4m3fn).view(torch.uint8).numpy()
