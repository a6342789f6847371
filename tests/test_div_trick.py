"""The CUDA quantiser divides with q0 = x*r, e = fma(-q0, s, x), q = fma(e, r, q0), r = RN(1/s)
(Markstein's correction) instead of an IEEE division per element.  This test emulates those three
float32 operations EXACTLY (rational arithmetic) and checks that the result equals RN(x / s) -- the
value vLLM's CUDA op computes -- for bf16 inputs and scales of the form RN(amax / 448)."""
from fractions import Fraction

import numpy as np


def rn32(v: Fraction) -> float:
    """Round a rational to the nearest float32 (ties to even), no overflow handling needed here."""
    if v == 0:
        return 0.0
    sgn = -1 if v < 0 else 1
    v = abs(v)
    e = v.numerator.bit_length() - v.denominator.bit_length()
    if Fraction(2) ** e > v:
        e -= 1
    e = max(e, -126)  # subnormals share the exponent of the smallest normal
    ulp = Fraction(2) ** (e - 23)
    q, rem = divmod(v, ulp)
    q = int(q)
    if rem * 2 > ulp or (rem * 2 == ulp and (q & 1)):
        q += 1
    return sgn * float(q * ulp)


def test_markstein_division_matches_ieee_for_bf16_inputs():
    rng = np.random.default_rng(0)
    n = 4000
    # bf16 values: random 16-bit patterns with moderate exponents, as float32
    def bf16(nv):
        mant = rng.integers(0, 128, nv).astype(np.uint32)
        exp = rng.integers(100, 140, nv).astype(np.uint32)
        return (((exp << 7) | mant) << 16).view(np.float32)

    amax = bf16(n)
    frac = rng.random(n).astype(np.float32)
    xs = (bf16(n) * 0 + amax * frac).astype(np.float32)
    xs = ((xs.view(np.uint32) >> 16) << 16).view(np.float32)  # truncate to a bf16 value <= amax
    xs[::7] = amax[::7]  # the element that defines amax
    bad = 0
    r448 = rn32(Fraction(1, 448))
    for a, x in zip(amax.tolist(), xs.tolist()):
        s = rn32(Fraction(a) / 448)
        # the kernel derives the scale the same way: q0 = a*r448, s = fma(fma(-q0, 448, a), r448, q0)
        s0 = rn32(Fraction(a) * Fraction(r448))
        es = rn32(Fraction(a) - Fraction(s0) * 448)
        assert rn32(Fraction(s0) + Fraction(es) * Fraction(r448)) == s
        r = rn32(1 / Fraction(s))
        q0 = rn32(Fraction(x) * Fraction(r))
        e = rn32(Fraction(x) - Fraction(q0) * Fraction(s))
        q = rn32(Fraction(q0) + Fraction(e) * Fraction(r))
        ref = rn32(Fraction(x) / Fraction(s))
        bad += q != ref
    assert bad == 0


def test_rn32_helper_matches_numpy():
    rng = np.random.default_rng(1)
    a = rng.standard_normal(500).astype(np.float32)
    b = (rng.standard_normal(500).astype(np.float32) + 3).astype(np.float32)
    for x, y in zip(a.tolist(), b.tolist()):
        assert rn32(Fraction(x) / Fraction(y)) == float(np.float32(x) / np.float32(y))
