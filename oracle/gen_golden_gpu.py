#!/usr/bin/env python
"""Generate tests/golden/vllm_fp8_quant_gpu.npz ON A GPU BOX (the build container has no GPU):

    gpurun -- 'python oracle/gen_golden_gpu.py gpurun_out/vllm_fp8_quant_gpu.npz'

The fixture holds the output of the reference's real quantiser -- vLLM's CUDA op
per_token_group_quant_8bit, reached through per_token_group_quant_fp8
(vllm/model_executor/layers/quantization/utils/fp8_utils.py:510-591) -- on seeded and adversarial
bf16 rows, with and without UE8M0 scales.  tests/test_oracle_path.py then pins the CPU oracle
(formula "div") against these bytes without needing a GPU.  Nothing here comes from the oracle."""
import sys

import numpy as np
import torch


def main(out):
    from vllm.model_executor.layers.quantization.utils.fp8_utils import per_token_group_quant_fp8

    H = 1024
    g = torch.Generator().manual_seed(20261106)
    x = (torch.randn(96, H, generator=g) * (torch.rand(96, 1, generator=g) * 4)).to(torch.bfloat16)
    adv = torch.zeros(8, H)
    adv[1, ::128] = 448.0
    adv[2] = 2.0**-20
    adv[3] = 3.0e38
    adv[4] = torch.linspace(-1, 1, H); adv[4, 7] = 1.75
    mids = [(1 + (m + 0.5) / 8) * 2.0**e for e in range(-6, 8) for m in range(8)]
    adv[5] = torch.tensor(mids).repeat(H // len(mids) + 1)[:H]; adv[5, ::128] = 448.0
    adv[6] = torch.randn(H, generator=g) * 1e-12
    adv[7] = torch.randn(H, generator=g) * 1e4
    advb = adv.to(torch.bfloat16)
    x = torch.cat([x, advb, -advb], 0).contiguous().cuda()
    save = {"x": x.cpu().view(torch.int16).numpy().view(np.uint16)}
    for ue in (False, True):
        q, s = per_token_group_quant_fp8(x, 128, use_ue8m0=ue)
        save[f"q_ue{int(ue)}"] = q.cpu().view(torch.uint8).numpy()
        save[f"s_ue{int(ue)}"] = s.contiguous().cpu().numpy()
    import vllm

    save["vllm_version"] = np.array(vllm.__version__)
    save["gpu"] = np.array(torch.cuda.get_device_name(0))
    np.savez_compressed(out, **save)
    print("wrote", out, {k: getattr(v, "shape", None) for k, v in save.items()})


if __name__ == "__main__":
    main(sys.argv[1])
