"""Worker for the multi-process tests (launched by torchrun; not collected by pytest).

  mode cpu : world_size ranks on the CPU over gloo -- host-side logic only: handle exchange and the
             rank / expert-ownership conventions of the oracle, checked rank by rank.
  mode gpu : one rank per GPU over CUDA IPC -- the real multi-GPU path, HT and LL, vs the oracle
             (tests/parity.py: the production all-phases calls; plus the recv-hook paths here).
  mode gpu_full : the same checks at the BASELINE.json shapes (4096 x 7168, 256/288 experts).
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "llm-d_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

from _util import bf16_bits, bits_to_bf16, make_rank_inputs, u8  # noqa: E402
from oracle import bits, combine, dispatch, fp8  # noqa: E402


def all_inputs(W, H, E, K, seed, Ts):
    ins = [make_rank_inputs(Ts[r], H, E, K, seed * 100 + r) for r in range(W)]
    return [i[0] for i in ins], [i[1] for i in ins], [i[2] for i in ins]


def run_cpu():
    dist.init_process_group("gloo")
    rank, W = dist.get_rank(), dist.get_world_size()
    from b200ep.buffer import exchange_handles

    blob = exchange_handles(dist.group.WORLD, bytes([rank]) * 64)
    assert blob == b"".join(bytes([r]) * 64 for r in range(W))
    # each rank contributes only ITS tokens; what it would receive must equal the single-process oracle
    H, E, K = 256, 16 * W, 4
    Ts = [13 + 5 * r for r in range(W)]
    xs, ids, ws = all_inputs(W, H, E, K, 3, Ts)
    mine = (bf16_bits(xs[rank]), ids[rank].numpy(), ws[rank].numpy())
    gathered = [None] * W
    dist.all_gather_object(gathered, mine)
    ref = dispatch.ht_dispatch([bf16_bits(x) for x in xs], [i.numpy() for i in ids], [w.numpy() for w in ws], E)
    got = dispatch.ht_dispatch([g[0] for g in gathered], [g[1] for g in gathered], [g[2] for g in gathered], E)
    assert np.array_equal(got[rank]["recv_x"], ref[rank]["recv_x"])
    assert np.array_equal(got[rank]["recv_topk_idx"], ref[rank]["recv_topk_idx"])
    # combine: every rank returns its received rows; the owner of the tokens sums them
    ys = [None] * W
    dist.all_gather_object(ys, got[rank]["recv_x"])
    out = combine.ht_combine(ys, got, Ts)[rank]
    ndst = np.zeros(Ts[rank])
    for d in range(W):
        pre = got[d]["rank_prefix"]
        ndst[got[d]["recv_src_idx"][pre[rank]:pre[rank + 1]]] += 1
    want = bits.f32_to_bf16(bits.bf16_to_f32(mine[0]) * ndst[:, None].astype(np.float32))
    # n identical bf16 rows summed in fp32 and rounded once == bf16(n * x) exactly for n <= 8... for n in {3,5,6,7} rounding may differ by <= 1 ulp
    diff = np.abs(bits.bf16_to_f32(out) - bits.bf16_to_f32(want))
    assert (diff <= np.abs(bits.bf16_to_f32(want)) * 2.0 ** -7).all()
    dist.barrier()
    print(f"CPU-MP-OK rank {rank}/{W}", flush=True)
    dist.destroy_process_group()


def _setup_gpu():
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    rank, W = dist.get_rank(), dist.get_world_size()
    cpu_group = dist.new_group(backend="gloo")

    def barrier():
        torch.cuda.synchronize()
        dist.barrier(group=cpu_group)

    return local, dev, rank, W, cpu_group, barrier


def run_gpu():
    """Small shapes, many variants: ragged and empty ranks, masked ids, both id dtypes, recv hooks,
    UE8M0 (fp32 and packed), the host-sync-free pair, 288 experts, and 20 back-to-back steps so that
    the call epochs advance well past both slots."""
    local, dev, rank, W, cpu_group, barrier = _setup_gpu()
    import b200ep
    import parity

    H, E, K = 1024, 64, 8
    buf = b200ep.Buffer(cpu_group, sym_bytes=512 << 20, device=local)
    # --- production calls (all phases per call): what vLLM and bench.py issue
    for it in range(20):
        Ts = [257 - 31 * r if (r + it) % 3 else 0 for r in range(W)]
        kind = ("uniform", "zipf", "one_rank")[it % 3]
        parity.check_ht(buf, cpu_group, W, rank, dev, Ts, H, E, K, seed=7 + it, dist=kind,
                        id_dtype=(torch.int32 if it % 4 == 1 else torch.int64), mask_frac=(0.2 if it % 5 == 2 else 0.0),
                        ue8m0=(it % 7 == 3), nosync=(it % 4 == 2))
        if it % 3 == 0:  # expert-grouped layout + combine from it, interleaved with the plain calls (shared slots, epochs)
            parity.check_ht_grouped(buf, cpu_group, W, rank, dev, Ts, H, E, K, seed=40 + it, dist=kind,
                                    mask_frac=(0.2 if it % 2 else 0.0))
    for E_ll in (64, 288):
        if E_ll % W:
            continue
        for graph in (False, True):
            case = parity.LlGraphCase(buf, cpu_group, W, rank, dev, 48 - rank, 2048, E_ll, K, 64, graph=graph,
                                      ue8m0=(E_ll == 288))
            for it in range(6):
                case.run(seed=300 + it, dist=("uniform", "zipf", "one_rank")[it % 3], mask_frac=(0.15 if it == 4 else 0.0),
                         barrier=barrier)
    bf = parity.LlGraphCase(buf, cpu_group, W, rank, dev, 33, 1024, 64, K, 64, use_fp8=False, graph=False)
    for it in range(3):
        bf.run(seed=330 + it, barrier=barrier)
    # --- recv-hook paths (split kernels), incl. packed UE8M0 scales, which only exist after the hook
    run_hook_paths(buf, rank, W, dev)
    torch.cuda.synchronize()
    dist.barrier()
    print(f"GPU-MP-OK rank {rank}/{W}", flush=True)
    buf.destroy()
    dist.destroy_process_group()


def run_hook_paths(buf, rank, W, dev):
    H, E, K, max_tok = 1024, 64, 8, 64
    for it in range(4):
        ue8m0 = bool(it & 1)
        Ts = [257 - 31 * r if (r + it) % 3 else 0 for r in range(W)]
        xs, ids, ws = all_inputs(W, H, E, K, 7 + it, Ts)
        Tl = [min(t, max_tok - r) for r, t in enumerate(Ts)]
        lx = [x[:n] for x, n in zip(xs, Tl)]
        li = [i[:n].contiguous() for i, n in zip(ids, Tl)]
        lw = [w[:n].contiguous() for w, n in zip(ws, Tl)]
        lq = [fp8.group_quant_fp8(bf16_bits(x), ue8m0=ue8m0) for x in lx]
        lref = dispatch.ll_dispatch([q for q, _ in lq], [i.numpy() for i in li], E, max_tok, scales=[s for _, s in lq])
        px, ps, cnt, lh, hook = buf.low_latency_dispatch(lx[rank].contiguous().to(dev), li[rank].to(dev), max_tok, E,
                                                         use_fp8=True, ue8m0=ue8m0, return_recv_hook=True)
        hook()
        torch.cuda.synchronize()
        lo = lref[rank]
        assert np.array_equal(cnt.cpu().numpy(), lo["recv_count"]), "ll counts"
        lys = []
        for d in range(W):
            y = np.zeros((E // W, W * max_tok, H), np.float32)
            for el in range(E // W):
                n = int(lref[d]["recv_count"][el])
                y[el, :n] = fp8.dequant_fp8(lref[d]["packed_recv_x"][el, :n], lref[d]["packed_recv_scales"][el, :n]) * (1 + el)
            lys.append(bits.f32_to_bf16(y.reshape(-1, H)).reshape(E // W, W * max_tok, H))
        pxh, psh = u8(px), ps.cpu().numpy()
        for el in range(E // W):
            n = int(lo["recv_count"][el])
            assert np.array_equal(pxh[el, :n], lo["packed_recv_x"][el, :n]), "ll fp8 bytes"
            assert np.array_equal(psh[el, :n], lo["packed_recv_scales"][el, :n]), "ll scales"
        lout, chook = buf.low_latency_combine(bits_to_bf16(lys[rank], dev), li[rank].to(dev), lw[rank].to(dev), lh,
                                              return_recv_hook=True)
        chook()
        torch.cuda.synchronize()
        lwant = combine.ll_combine(lys, [i.numpy() for i in li], [w.numpy() for w in lw], lref, E)[rank]
        assert np.array_equal(bf16_bits(lout), lwant), "ll combine"
        dist.barrier()


def run_gpu_full():
    """BASELINE.json shapes through the production calls: HT 4096 tokens/rank, hidden 7168, 256 experts,
    top-8 (uniform / Zipf / all-to-one-rank routing, synchronous and host-sync-free); LL 128 tokens/rank,
    max 256, 256 and 288 experts, captured in a CUDA graph and replayed with fresh inputs."""
    local, dev, rank, W, cpu_group, barrier = _setup_gpu()
    import b200ep
    import parity

    T, H, E, K = 4096, 7168, 256, 8
    need = 2 * (W * T * (H + H // 32 + K * 12 + 4) + T * W * 2 * H) + (64 << 20)
    ll_need = max(b200ep.Buffer.ll_size_hint(256, H, W, 288 if 288 % W == 0 else E), b200ep.Buffer.ll_size_hint(256, H, W, E))
    buf = b200ep.Buffer(cpu_group, sym_bytes=int(max(need, ll_need) * 1.1), device=local)
    out = torch.empty((T, H), dtype=torch.bfloat16, device=dev)
    for it, kind in enumerate(("uniform", "zipf", "one_rank")):
        parity.check_ht(buf, cpu_group, W, rank, dev, [T] * W, H, E, K, seed=50 + it, dist=kind, combine_out=out)
        parity.check_ht(buf, cpu_group, W, rank, dev, [T] * W, H, E, K, seed=60 + it, dist=kind, nosync=True, combine_out=out)
    parity.check_ht_grouped(buf, cpu_group, W, rank, dev, [T] * W, H, E, K, seed=70)
    for E_ll in (256, 288):
        if E_ll % W:
            continue
        case = parity.LlGraphCase(buf, cpu_group, W, rank, dev, 128, H, E_ll, K, 256, graph=True)
        for it, kind in enumerate(("uniform", "zipf", "one_rank", "uniform")):
            case.run(seed=80 + it, dist=kind, barrier=barrier)
    torch.cuda.synchronize()
    dist.barrier()
    print(f"GPU-MP-FULL-OK rank {rank}/{W}", flush=True)
    buf.destroy()
    dist.destroy_process_group()


if __name__ == "__main__":
    {"cpu": run_cpu, "gpu": run_gpu, "gpu_full": run_gpu_full}[sys.argv[1]]()
