#!/usr/bin/env python
"""Generate the CPU golden fixtures under tests/golden/ (run in the build
container; the GPU box never runs this).

Each fixture is produced by IMPORTING the installed vLLM 0.22.0 (the reference
pins v0.23.0, ``docker/common-versions:5-6``) and running the real function on
seeded inputs -- none of the outputs below come from the oracle:

  cpu_moe_granite.npz   ``cpu_fused_moe_torch`` (cpu_fused_moe.py:428-481) with
                        routing by ``select_experts`` (:117-161), on the
                        BASELINE.json config #1 shape (32 experts, top-8,
                        hidden 1024).  The expert MLPs are replaced by
                        per-expert scalings so that every output row depends
                        on which expert processed it.
  batched_prepare.npz   ``BatchedPrepareAndFinalize.prepare`` (batched.py:55-152),
                        unquantised, for rank 0..W-1 of W in {1, 2, 4}: the
                        BatchedExperts layout the low-latency path must emit.
  casts.npz             torch's CPU ``float8_e4m3fn`` / ``bfloat16`` casts on an
                        adversarial float32 set (rounding midpoints, subnormals,
                        saturation, signed zeros).

Usage:  python oracle/gen_golden.py
"""

from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden")


def bf16_bits(t: torch.Tensor) -> np.ndarray:
    return t.contiguous().view(torch.int16).numpy().view(np.uint16).copy()


def gen_cpu_moe():
    import vllm.model_executor.layers.fused_moe.cpu_fused_moe as cm
    from vllm.model_executor.layers.fused_moe.activation import MoEActivation

    E, K, H, T = 32, 8, 1024, 96
    g = torch.Generator().manual_seed(20260101)
    x = (torch.randn(T, H, generator=g) * (torch.rand(T, 1, generator=g) * 4)).to(torch.bfloat16)
    logits = torch.randn(T, E, generator=g)
    w, ids = cm.select_experts(x, logits, K, use_grouped_topk=False, renormalize=True)

    coef = (1.0 + torch.arange(E, dtype=torch.float32) / E).to(torch.bfloat16)

    class _Layer:  # what _CPU_MOE_LAYER_CACHE hands to cpu_fused_moe_torch
        gate_up_linear = [lambda r: r for _ in range(E)]
        down_linear = [(lambda r, i=i: r * coef[i]) for i in range(E)]

    layer = _Layer()
    cm._CPU_MOE_LAYER_CACHE[12345] = lambda: layer
    saved = cm._CPU_MOE_ACT_FN[MoEActivation.SILU]
    cm._CPU_MOE_ACT_FN[MoEActivation.SILU] = lambda r: r  # identity activation
    try:
        out = torch.empty_like(x)
        cm.cpu_fused_moe_torch(12345, out, x, w, ids, "silu", global_num_experts=E)
    finally:
        cm._CPU_MOE_ACT_FN[MoEActivation.SILU] = saved
        del cm._CPU_MOE_LAYER_CACHE[12345]
    np.savez_compressed(
        os.path.join(OUT, "cpu_moe_granite.npz"),
        x=bf16_bits(x),
        logits=logits.numpy(),
        topk_ids=ids.numpy(),
        topk_weights=w.numpy(),
        coef=bf16_bits(coef),
        out=bf16_bits(out),
    )


def gen_batched():
    from vllm.model_executor.layers.fused_moe.config import FUSED_MOE_UNQUANTIZED_CONFIG
    from vllm.model_executor.layers.fused_moe.prepare_finalize.batched import BatchedPrepareAndFinalize

    E, K, H, T, max_tok = 16, 4, 256, 40, 48
    g = torch.Generator().manual_seed(7)
    x = torch.randn(T, H, generator=g).to(torch.bfloat16)
    ids = torch.stack([torch.randperm(E, generator=g)[:K] for _ in range(T)]).to(torch.int32)
    w = torch.rand(T, K, generator=g)
    save = dict(x=bf16_bits(x), topk_ids=ids.numpy(), topk_weights=w.numpy(), max_tok=max_tok, E=E)
    for W in (1, 2, 4):
        for r in range(W):
            pf = BatchedPrepareAndFinalize(max_tok, E // W, W, r)
            b_a1, _, meta, _, _ = pf.prepare(x.clone(), w, ids, E, None, False, FUSED_MOE_UNQUANTIZED_CONFIG)
            save[f"w{W}_r{r}_x"] = bf16_bits(b_a1)
            save[f"w{W}_r{r}_cnt"] = meta.expert_num_tokens.numpy()
    np.savez_compressed(os.path.join(OUT, "batched_prepare.npz"), **save)


def gen_casts():
    rng = np.random.default_rng(3)
    vals = [0.0, -0.0, 2.0**-10, 2.0**-9, 1.5 * 2.0**-9, 2.5 * 2.0**-9, 2.0**-6, 447.9, 448.0, 449.0, 463.9,
            464.0, 465.0, 480.0, 1e6, -1e6, float("inf"), -float("inf"), 2.0**-7 + 2.0**-10]
    # every e4m3 midpoint, just below and just above it
    for e in range(-6, 9):
        for m in range(8):
            lo = (1 + m / 8) * 2.0**e
            mid = lo + 2.0 ** (e - 4)
            vals += [lo, mid, np.nextafter(np.float32(mid), np.float32(0)), np.nextafter(np.float32(mid), np.float32(1e9))]
    for m in range(8):
        mid = (m + 0.5) * 2.0**-9
        vals += [mid, np.nextafter(np.float32(mid), np.float32(0)), np.nextafter(np.float32(mid), np.float32(1))]
    v = np.array(vals, np.float32)
    v = np.concatenate([v, -v, (rng.standard_normal(4096) * 100).astype(np.float32),
                        (rng.standard_normal(4096) * 0.01).astype(np.float32)])
    t = torch.from_numpy(v)
    e4 = t.clamp(-448.0, 448.0).to(torch.float8_e4m3fn).view(torch.uint8).numpy()
    b16 = bf16_bits(t.to(torch.bfloat16))
    np.savez_compressed(os.path.join(OUT, "casts.npz"), f32=v, e4m3_sat=e4, bf16=b16)


def main():
    os.makedirs(OUT, exist_ok=True)
    gen_casts()
    gen_batched()
    gen_cpu_moe()
    print("wrote", sorted(os.listdir(OUT)))


if __name__ == "__main__":
    sys.exit(main())
