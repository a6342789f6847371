"""Worker for tests/test_multiproc.py::test_gpu_vllm_managers_world (launched by torchrun, one rank per GPU).

vLLM itself builds everything here: ``ParallelConfig(data_parallel_size=W, enable_expert_parallel=True,
all2all_backend="deepep_*")`` + ``initialize_model_parallel`` create the EP group, whose CudaCommunicator
constructs vLLM's own ``DeepEPHTAll2AllManager`` / ``DeepEPLLAll2AllManager``
(vllm/distributed/device_communicators/cuda_communicator.py:121-175, all2all.py:143-325); ``get_handle`` then
builds a ``deep_ep.Buffer`` -- which is llm-d_b200/deep_ep, the b200ep shim, because that directory is on the
path (what the llm-d manifests get with PYTHONPATH set: decode.yaml:119 / prefill.yaml:91 keep their flags).
vLLM's own prepare/finalize classes are then driven on that handle at world size W, with TWO micro-batches in
flight on two streams (two layers' worth of dispatches before the first combine: the Buffer-level picture of
--enable-dbo), and every rank's results are compared with the CPU oracle.

    torchrun --nproc-per-node W tests/mp_vllm_worker.py deepep_high_throughput|deepep_low_latency
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "llm-d_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

from _util import bf16_bits, bits_to_bf16, make_rank_inputs, u8  # noqa: E402
from oracle import bits, combine, dispatch, fp8, layout  # noqa: E402


def gather(group, obj, W):
    import torch.distributed as dist

    out = [None] * W
    dist.all_gather_object(out, obj, group=group)
    return out


def main():
    backend = sys.argv[1]
    rank, W, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)

    import deep_ep
    from vllm.config import ParallelConfig, VllmConfig, set_current_vllm_config
    from vllm.distributed.parallel_state import (get_ep_group, init_distributed_environment,
                                                 initialize_model_parallel)

    assert deep_ep.__file__.startswith(os.path.join(ROOT, "llm-d_b200")), "deep_ep must resolve to the b200ep shim"
    pc = ParallelConfig(data_parallel_size=W, enable_expert_parallel=True, all2all_backend=backend,
                        distributed_executor_backend="external_launcher")
    vc = VllmConfig(parallel_config=pc)
    with set_current_vllm_config(vc):
        init_distributed_environment(world_size=W, rank=rank, distributed_init_method="env://", local_rank=local,
                                     backend="nccl")
        initialize_model_parallel(1, 1)
        ep = get_ep_group()
        mgr = ep.device_communicator.all2all_manager
        want_cls = {"deepep_high_throughput": "DeepEPHTAll2AllManager", "deepep_low_latency": "DeepEPLLAll2AllManager"}[backend]
        assert type(mgr).__name__ == want_cls, f"vLLM built {type(mgr).__name__}, expected {want_cls}"
        assert mgr.world_size == W and mgr.rank == rank and not mgr.internode
        if backend == "deepep_high_throughput":
            run_ht(mgr, ep.cpu_group, rank, W, dev)
        else:
            run_ll(mgr, ep.cpu_group, rank, W, dev)
        torch.cuda.synchronize()
        torch.distributed.barrier(group=ep.cpu_group)
        mgr.destroy()
    print(f"VLLM-MGR-OK {backend} rank {rank}/{W}", flush=True)


def quant_config():
    from vllm.model_executor.layers.fused_moe.config import fp8_w8a8_moe_quant_config

    one = torch.ones(1, device="cuda")
    return fp8_w8a8_moe_quant_config(w1_scale=one, w2_scale=one, block_shape=[128, 128])


def run_ht(mgr, cpu_group, rank, W, dev):
    import deep_ep
    from vllm.model_executor.layers.fused_moe.prepare_finalize.deepep_ht import DeepEPHTPrepareAndFinalize
    from vllm.model_executor.layers.fused_moe.topk_weight_and_reduce import TopKWeightAndReduceNoOP

    H, E, K = 2048, 64, 8
    el = E // W
    handle = mgr.get_handle({})
    assert isinstance(handle, deep_ep.Buffer) and mgr.get_handle({}) is handle  # cached by kwargs
    assert handle.group_size == W and handle.rank == rank
    qc = quant_config()
    # two layers' prepare/finalize objects on the ONE buffer, each on a stream of its own
    pfs = [DeepEPHTPrepareAndFinalize(handle, num_dispatchers=W, dp_size=W, rank_expert_offset=rank * el) for _ in range(2)]
    streams = [torch.cuda.Stream(device=dev) for _ in range(2)]
    for it in range(3):
        cases = []
        for u in range(2):  # micro-batch u
            Ts = [150 - 17 * r + 5 * u if (r + it + u) % 4 else 3 for r in range(W)]  # ragged; never empty (vLLM's quantiser)
            x, ids, w = make_rank_inputs(Ts[rank], H, E, K, 1000 * it + 10 * u + rank, dist=("uniform", "zipf")[u], world=W)
            cases.append((Ts, x, ids, w))
        prepared = []
        torch.cuda.synchronize()
        for u, (Ts, x, ids, w) in enumerate(cases):  # both dispatches before any combine
            with torch.cuda.stream(streams[u]):
                prepared.append(pfs[u].prepare(x.to(dev), w.to(dev), ids.to(dev), E, None, False, qc, False))
        for u, (Ts, x, ids, w) in enumerate(cases):
            a1q, a1s, meta, oids, ow = prepared[u]
            streams[u].synchronize()
            # the quantiser vLLM ran before the dispatch decides the UE8M0 mode; the oracle must match that mode
            modes = {False: fp8.group_quant_fp8(bf16_bits(x), ue8m0=False), True: fp8.group_quant_fp8(bf16_bits(x), ue8m0=True)}
            allr = gather(cpu_group, (modes, ids.numpy(), w.numpy()), W)
            got_s = a1s.contiguous().cpu().numpy() if a1s is not None else None
            ok = False
            for ue in (False, True):
                ref = dispatch.ht_dispatch([a[0][ue][0] for a in allr], [a[1] for a in allr], [a[2] for a in allr], E,
                                           scales=[a[0][ue][1] for a in allr], dests=[rank])[rank]
                if got_s is not None and got_s.shape == ref["recv_x_scales"].shape and np.array_equal(got_s, ref["recv_x_scales"]):
                    ok = True
                    break
            assert ok, f"rank {rank}: received scales match neither quantiser mode of the oracle"
            M = ref["num_recv_tokens"]
            assert a1q.shape[0] == M and np.array_equal(u8(a1q), ref["recv_x"]), "received fp8 rows"
            assert np.array_equal(meta.expert_num_tokens.cpu().numpy(), ref["num_recv_tokens_per_expert"])
            assert meta.expert_num_tokens_cpu.tolist() == ref["num_recv_tokens_per_expert"].tolist()
            lid = ref["recv_topk_idx"]
            sentinel = E - 1 if rank * el == 0 else 0  # deepep_ht.py:204-220 maps -1 to an expert that is NOT local
            want_ids = np.where(lid == -1, sentinel, lid + rank * el)
            assert np.array_equal(oids.cpu().numpy(), want_ids), "expert_topk_ids"
            assert np.array_equal(ow.cpu().numpy(), ref["recv_topk_weights"]), "expert_topk_weights"
            y_np = bits.f32_to_bf16(fp8.dequant_fp8(ref["recv_x"], ref["recv_x_scales"]) * np.float32(1 + 0.25 * rank))
            out = torch.full((cases[u][0][rank], H), float("nan"), dtype=torch.bfloat16, device=dev)
            with torch.cuda.stream(streams[u]):
                y = bits_to_bf16(y_np, dev).view(M, H)
                pfs[u].finalize(out, y, ow, oids, False, TopKWeightAndReduceNoOP())
            streams[u].synchronize()
            q, s = modes[ue]
            in_rank = layout.dispatch_layout(ids.numpy(), E, W)[2]
            deq = fp8.dequant_fp8(q, s)
            parts = [(np.nonzero(in_rank[:, d])[0], bits.f32_to_bf16(deq[np.nonzero(in_rank[:, d])[0]] * np.float32(1 + 0.25 * d)))
                     for d in range(W)]
            want = combine.ht_combine_source(cases[u][0][rank], H, parts)
            assert np.array_equal(bf16_bits(out), want), f"rank {rank}: finalize output (iteration {it}, micro-batch {u})"


def run_ll(mgr, cpu_group, rank, W, dev):
    import deep_ep
    from vllm.model_executor.layers.fused_moe.prepare_finalize.deepep_ll import DeepEPLLPrepareAndFinalize
    from vllm.model_executor.layers.fused_moe.topk_weight_and_reduce import TopKWeightAndReduceDelegate

    H, E, K, max_tok = 2048, 64, 8, 64
    el = E // W
    kw = dict(max_num_tokens_per_dp_rank=max_tok, token_hidden_size=H, num_ep_ranks=W, num_global_experts=E, num_local_experts=el)
    handle = mgr.get_handle(kw)
    assert isinstance(handle, deep_ep.Buffer) and handle.low_latency_mode and mgr.get_handle(dict(kw)) is handle
    qc = quant_config()
    for packed in (False, True):
        pfs = [DeepEPLLPrepareAndFinalize(handle, max_tokens_per_rank=max_tok, num_dispatchers=W, use_fp8_dispatch=True)
               for _ in range(2)]
        for pf in pfs:
            pf.use_ue8m0_dispatch = packed  # what post_init_setup does for experts that take packed UE8M0 scales
        streams = [torch.cuda.Stream(device=dev) for _ in range(2)]
        for it in range(3):
            cases = []
            for u in range(2):
                T = max_tok - 3 * rank - u - it
                cases.append(make_rank_inputs(T, H, E, K, 2000 * it + 10 * u + rank, dist=("uniform", "zipf")[u], world=W))
            prepared = []
            torch.cuda.synchronize()
            for u, (x, ids, w) in enumerate(cases):
                with torch.cuda.stream(streams[u]):
                    prepared.append(pfs[u].prepare(x.to(dev), w.to(dev), ids.to(dev), E, None, False, qc))
            for u, (x, ids, w) in enumerate(cases):
                ex, es, meta, _, _ = prepared[u]
                streams[u].synchronize()
                q, s = fp8.group_quant_fp8(bf16_bits(x), ue8m0=packed)
                allr = gather(cpu_group, (q, s, ids.numpy()), W)
                ref = dispatch.ll_dispatch([a[0] for a in allr], [a[2] for a in allr], E, max_tok,
                                           scales=[a[1] for a in allr], dests=[rank])[rank]
                assert np.array_equal(meta.expert_num_tokens.cpu().numpy(), ref["recv_count"]), "expert_num_tokens"
                exh, esh = u8(ex), es.cpu().numpy()
                y = np.zeros((el, W * max_tok, H), np.uint16)
                for e in range(el):
                    n = int(ref["recv_count"][e])
                    assert np.array_equal(exh[e, :n], ref["packed_recv_x"][e, :n]), "payload"
                    if packed:
                        assert es.dtype == torch.int32 and np.array_equal(esh[e, :n], fp8.pack_ue8m0(ref["packed_recv_scales"][e, :n]))
                    else:
                        assert np.array_equal(esh[e, :n], ref["packed_recv_scales"][e, :n]), "scales"
                    y[e, :n] = bits.f32_to_bf16(fp8.dequant_fp8(ref["packed_recv_x"][e, :n], ref["packed_recv_scales"][e, :n])
                                                * np.float32(1 + 0.125 * rank + 0.5 * e))
                out = torch.full((x.shape[0], H), float("nan"), dtype=torch.bfloat16, device=dev)
                with torch.cuda.stream(streams[u]):
                    yd = bits_to_bf16(y.reshape(-1, H), dev).view(el, W * max_tok, H)
                    pfs[u].finalize(out, yd, w.to(dev), ids.to(dev), False, TopKWeightAndReduceDelegate())
                streams[u].synchronize()
                mine = fp8.dequant_fp8(q, s)
                idn = ids.numpy().astype(np.int64)
                rows = np.zeros((x.shape[0], K, H), np.uint16)
                for k in range(K):
                    f = (1.0 + 0.125 * (idn[:, k] // el) + 0.5 * (idn[:, k] % el)).astype(np.float32)
                    rows[:, k] = bits.f32_to_bf16(mine * f[:, None])
                want = combine.ll_combine_source(idn, w.numpy(), rows, E)
                assert np.array_equal(bf16_bits(out), want), f"rank {rank}: LL finalize (packed={packed}, iteration {it}, micro-batch {u})"


if __name__ == "__main__":
    main()
