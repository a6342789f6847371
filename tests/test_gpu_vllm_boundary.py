"""Boundary test: vLLM's OWN prepare/finalize classes (DeepEPHTPrepareAndFinalize,
DeepEPLLPrepareAndFinalize -- the code llm-d's --all2all-backend deepep_* flags select) running
unmodified on the deep_ep-compatible shim, plus the fused-quantise subclass of Route B.
One rank (vLLM itself never builds an all2all manager at dp_size 1, so this drives the classes
directly); results are compared with the CPU oracle."""
import os
import sys

import numpy as np
import pytest
import torch

from _util import bf16_bits, make_rank_inputs, u8
from oracle import bits, fp8

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def group():
    import torch.distributed as dist

    if not dist.is_initialized():
        dist.init_process_group("gloo", rank=0, world_size=1, init_method="tcp://127.0.0.1:29541")
    yield dist.group.WORLD


def _vllm_bits():
    try:
        import deep_ep  # noqa: F401  (the shim: tests/conftest.py puts llm-d_b200 on sys.path)
        import vllm.model_executor.layers.fused_moe.modular_kernel as mk  # noqa: F401
        from vllm.model_executor.layers.fused_moe.config import fp8_w8a8_moe_quant_config
        from vllm.model_executor.layers.fused_moe.prepare_finalize.deepep_ht import DeepEPHTPrepareAndFinalize
        from vllm.model_executor.layers.fused_moe.prepare_finalize.deepep_ll import DeepEPLLPrepareAndFinalize
        from vllm.model_executor.layers.fused_moe.topk_weight_and_reduce import (TopKWeightAndReduceDelegate,
                                                                                 TopKWeightAndReduceNoOP)
    except Exception as e:  # pragma: no cover
        pytest.skip(f"vLLM not importable: {e}")
    return (fp8_w8a8_moe_quant_config, DeepEPHTPrepareAndFinalize, DeepEPLLPrepareAndFinalize,
            TopKWeightAndReduceDelegate, TopKWeightAndReduceNoOP)


@pytest.mark.parametrize("fused", [False, True])
def test_vllm_deepep_ht_prepare_finalize_on_shim(group, fused):
    import deep_ep

    mkq, HT, _, _, NoOP = _vllm_bits()
    torch.cuda.set_device(0)
    T, H, E, K = 333, 2048, 64, 8
    x, ids, w = make_rank_inputs(T, H, E, K, seed=11)
    buf = deep_ep.Buffer(group=group, num_nvl_bytes=1 << 30, num_rdma_bytes=0, low_latency_mode=False,
                         num_qps_per_rank=1, explicitly_destroy=True)
    try:
        if fused:
            from b200ep.vllm_backend import get_classes

            HT = get_classes()["B200EPHTPrepareAndFinalize"]
        pf = HT(buf, num_dispatchers=1, dp_size=1, rank_expert_offset=0)
        assert pf.topk_indices_dtype() == torch.int64 and pf.output_is_reduced()
        qc = mkq(w1_scale=torch.ones(1, device=DEV), w2_scale=torch.ones(1, device=DEV), block_shape=[128, 128])
        a1q, a1s, meta, oids, ow = pf.prepare(x.to(DEV), w.to(DEV), ids.to(DEV), E, None, False, qc, False)
        torch.cuda.synchronize()
        # ue8m0 is decided by vLLM (is_deep_gemm_e8m0_used); accept either, but it must match the oracle of that mode
        oq0, os0 = fp8.group_quant_fp8(bf16_bits(x), formula="div", ue8m0=False)
        oq1, os1 = fp8.group_quant_fp8(bf16_bits(x), formula="div", ue8m0=True)
        got_s = a1s.contiguous().cpu().numpy()
        assert np.array_equal(got_s, os0) or np.array_equal(got_s, os1)
        oq = oq0 if np.array_equal(got_s, os0) else oq1
        assert a1q.shape == (T, H) and np.array_equal(u8(a1q), oq)  # one rank: every token, ascending order
        assert np.array_equal(meta.expert_num_tokens.cpu().numpy(), np.bincount(ids.numpy().ravel(), minlength=E))
        assert meta.expert_num_tokens_cpu.tolist() == meta.expert_num_tokens.cpu().tolist()
        assert torch.equal(oids.cpu(), ids) and torch.equal(ow.cpu(), w)
        # finalize: experts' (already weighted and reduced) rows come back to their tokens
        y = torch.randn(T, H, device=DEV).to(torch.bfloat16)
        out = torch.zeros(T, H, dtype=torch.bfloat16, device=DEV)
        pf.finalize(out, y, w.to(DEV), ids.to(DEV), False, NoOP())
        torch.cuda.synchronize()
        assert torch.equal(out, y)
    finally:
        buf.destroy()


def test_vllm_deepep_ll_prepare_finalize_on_shim(group):
    import deep_ep

    mkq, _, LL, Delegate, _ = _vllm_bits()
    torch.cuda.set_device(0)
    T, H, E, K, max_tok = 100, 2048, 32, 8, 128
    x, ids, w = make_rank_inputs(T, H, E, K, seed=12)
    nbytes = deep_ep.Buffer.get_low_latency_rdma_size_hint(max_tok, H, 1, E)
    buf = deep_ep.Buffer(group=group, num_nvl_bytes=0, num_rdma_bytes=nbytes, low_latency_mode=True,
                         num_qps_per_rank=E, allow_nvlink_for_low_latency_mode=True, allow_mnnvl=False,
                         explicitly_destroy=True)
    try:
        pf = LL(buf, max_tokens_per_rank=max_tok, num_dispatchers=1, use_fp8_dispatch=True)
        assert pf.max_num_tokens_per_rank() == max_tok
        qc = mkq(w1_scale=torch.ones(1, device=DEV), w2_scale=torch.ones(1, device=DEV), block_shape=[128, 128])
        ex, es, meta, _, _ = pf.prepare(x.to(DEV), w.to(DEV), ids.to(DEV), E, None, False, qc)
        torch.cuda.synchronize()
        cnt = meta.expert_num_tokens.cpu().numpy()
        assert np.array_equal(cnt, np.bincount(ids.numpy().ravel(), minlength=E))
        oq, os_ = fp8.group_quant_fp8(bf16_bits(x), formula="div")
        exh, esh = u8(ex), es.cpu().numpy()
        y = torch.zeros(E, max_tok, H, dtype=torch.bfloat16, device=DEV)
        for e in range(E):
            rows = np.nonzero((ids.numpy() == e).any(1))[0]  # ascending tokens == batched.py:115-122
            assert np.array_equal(exh[e, : len(rows)], oq[rows])
            assert np.array_equal(esh[e, : len(rows)], os_[rows])
            deq = fp8.dequant_fp8(oq[rows], os_[rows])
            y[e, : len(rows)] = torch.from_numpy(deq).to(torch.bfloat16).to(DEV)
        out = torch.zeros(T, H, dtype=torch.bfloat16, device=DEV)
        pf.finalize(out, y, w.to(DEV), ids.to(DEV), False, Delegate())
        torch.cuda.synchronize()
        ref = x.float()  # weights sum to 1: sum_k w_k * dequant(x) ~ x
        assert (out.cpu().float() - ref).abs().max().item() <= 0.08 * ref.abs().max().item()
    finally:
        buf.destroy()


def test_host_sync_free_prepare_finalize_on_shim(group):
    """Route B, f2: B200EPHTNoSyncPrepareAndFinalize never waits on the host; counts come back on the device
    (expert_num_tokens_cpu is None) and rows beyond the received count are neutralised."""
    import deep_ep
    from b200ep.vllm_backend import get_classes

    mkq, _, _, _, NoOP = _vllm_bits()
    torch.cuda.set_device(0)
    T, H, E, K, cap = 200, 2048, 64, 8, 256
    x, ids, w = make_rank_inputs(T, H, E, K, seed=21, mask_frac=0.2)
    ids[5] = -1  # a token without any expert: fewer rows than tokens
    buf = deep_ep.Buffer(group=group, num_nvl_bytes=1 << 30, num_rdma_bytes=0, low_latency_mode=False,
                         num_qps_per_rank=1, explicitly_destroy=True)
    try:
        pf = get_classes()["B200EPHTNoSyncPrepareAndFinalize"](buf, num_dispatchers=1, dp_size=1, rank_expert_offset=0,
                                                              max_tokens_per_rank=cap)
        qc = mkq(w1_scale=torch.ones(1, device=DEV), w2_scale=torch.ones(1, device=DEV), block_shape=[128, 128])
        a1q, a1s, meta, oids, ow = pf.prepare(x.to(DEV), w.to(DEV), ids.to(DEV), E, None, False, qc, False)
        torch.cuda.synchronize()
        assert meta.expert_num_tokens_cpu is None and a1q.shape == (cap, H)
        valid = (ids >= 0).any(1).numpy()
        rows = np.nonzero(valid)[0]
        M = len(rows)
        assert np.array_equal(meta.expert_num_tokens.cpu().numpy(), np.bincount(ids.numpy()[ids.numpy() >= 0], minlength=E))
        got_s = a1s.cpu().numpy()[:M]
        oq0, os0 = fp8.group_quant_fp8(bf16_bits(x), ue8m0=False)
        oq1, os1 = fp8.group_quant_fp8(bf16_bits(x), ue8m0=True)
        oq, osc = (oq0, os0) if np.array_equal(got_s, os0[rows]) else (oq1, os1)
        assert np.array_equal(got_s, osc[rows]) and np.array_equal(u8(a1q)[:M], oq[rows])
        want_ids = np.where(ids.numpy()[rows] >= 0, ids.numpy()[rows], E - 1)
        assert np.array_equal(oids.cpu().numpy()[:M], want_ids)
        assert (oids.cpu().numpy()[M:] == E - 1).all() and (ow.cpu().numpy()[M:] == 0).all()
        y = torch.full((cap, H), float("nan"), dtype=torch.bfloat16, device=DEV)
        y[:M] = torch.randn(M, H, device=DEV).to(torch.bfloat16)
        out = torch.full((T, H), float("nan"), dtype=torch.bfloat16, device=DEV)
        pf.finalize(out, y, ow, oids, False, NoOP())
        torch.cuda.synchronize()
        assert torch.equal(out[torch.from_numpy(rows).to(DEV)], y[:M])
        assert bool((out[torch.from_numpy(~valid).to(DEV)] == 0).all())
    finally:
        buf.destroy()
