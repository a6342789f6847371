"""Shared helpers for the GPU parity tests (seeded synthetic inputs)."""
import numpy as np
import torch


def bf16_bits(t: torch.Tensor) -> np.ndarray:
    return t.detach().cpu().contiguous().view(torch.int16).numpy().view(np.uint16)


def u8(t: torch.Tensor) -> np.ndarray:
    return t.detach().cpu().contiguous().view(torch.uint8).numpy()


def bits_to_bf16(a: np.ndarray, device) -> torch.Tensor:
    return torch.from_numpy(a.view(np.int16).copy()).view(torch.bfloat16).to(device)


def make_rank_inputs(T, H, E, K, seed, dist="uniform", world=1, id_dtype=torch.int64, mask_frac=0.0):
    """x ~ N(0,1)*U(0,4) per row (bf16), router logits N(0,1) -> top-k ids, softmax weights (fp32)."""
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(T, H, generator=g) * (torch.rand(T, 1, generator=g) * 4)).to(torch.bfloat16)
    logits = torch.randn(T, E, generator=g)
    if dist == "one_rank":  # every token routes to the last rank only
        logits[:, : E - E // world] = -1e9
    elif dist == "zipf":
        logits += torch.log(1.0 / torch.arange(1, E + 1, dtype=torch.float32) ** 1.2)[None, :]
    vals, ids = torch.topk(logits, K, dim=-1, sorted=False)
    w = torch.softmax(vals, dim=-1).float()
    ids = ids.to(id_dtype)
    if mask_frac > 0:
        m = torch.rand(T, K, generator=g) < mask_frac
        ids[m] = -1
    return x.contiguous(), ids.contiguous(), w.contiguous()
