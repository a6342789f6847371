"""Worker for the multi-process tests (launched by torchrun; not collected by pytest).

  mode cpu : world_size ranks on the CPU over gloo -- host-side logic only: handle exchange and the
             rank / expert-ownership conventions of the oracle, checked rank by rank.
  mode gpu : one rank per GPU over CUDA IPC -- the real multi-GPU path, HT and LL, vs the oracle.
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


def run_gpu():
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    rank, W = dist.get_rank(), dist.get_world_size()
    cpu_group = dist.new_group(backend="gloo")
    import b200ep

    H, E, K = 1024, 64, 8
    buf = b200ep.Buffer(cpu_group, sym_bytes=512 << 20, device=local)
    for it in range(4):
        Ts = [257 - 31 * r if (r + it) % 3 else 0 for r in range(W)]
        xs, ids, ws = all_inputs(W, H, E, K, 7 + it, Ts)
        xb = [bf16_bits(x) for x in xs]
        qs = [fp8.group_quant_fp8(b) for b in xb]
        ref = dispatch.ht_dispatch([q for q, _ in qs], [i.numpy() for i in ids], [w.numpy() for w in ws], E,
                                   scales=[s for _, s in qs])
        rx, rs, ridx, rw, per_expert, h = buf.dispatch(xs[rank].to(dev), ids[rank].to(dev), ws[rank].to(dev), E,
                                                       quantize=True)
        o = ref[rank]
        assert h.num_recv == o["num_recv_tokens"] and per_expert == list(o["num_recv_tokens_per_expert"])
        assert np.array_equal(u8(rx), o["recv_x"]), "fp8 bytes"
        assert np.array_equal(rs.cpu().numpy(), o["recv_x_scales"]), "scales"
        assert np.array_equal(ridx.cpu().numpy(), o["recv_topk_idx"]), "topk idx"
        assert np.array_equal(rw.cpu().numpy(), o["recv_topk_weights"]), "topk weights"
        ys = [bits.f32_to_bf16(fp8.dequant_fp8(r_["recv_x"], r_["recv_x_scales"]) * (1 + 0.25 * d)) for d, r_ in enumerate(ref)]
        out = buf.combine(bits_to_bf16(ys[rank], dev), h)
        want = combine.ht_combine(ys, ref, Ts)[rank]
        assert np.array_equal(bf16_bits(out), want), "combine"
        # low latency
        max_tok = 64
        Tl = [min(t, max_tok - r) for r, t in enumerate(Ts)]
        lx = [x[:n] for x, n in zip(xs, Tl)]
        li = [i[:n].contiguous() for i, n in zip(ids, Tl)]
        lw = [w[:n].contiguous() for w, n in zip(ws, Tl)]
        lq = [fp8.group_quant_fp8(bf16_bits(x)) for x in lx]
        lref = dispatch.ll_dispatch([q for q, _ in lq], [i.numpy() for i in li], E, max_tok, scales=[s for _, s in lq])
        px, ps, cnt, lh, hook = buf.low_latency_dispatch(lx[rank].contiguous().to(dev), li[rank].to(dev), max_tok, E,
                                                         use_fp8=True, return_recv_hook=True)
        hook()
        lo = lref[rank]
        assert np.array_equal(cnt.cpu().numpy(), lo["recv_count"]), "ll counts"
        lys = []
        for d in range(W):
            y = np.zeros((E // W, W * max_tok, H), np.float32)
            for el in range(E // W):
                n = int(lref[d]["recv_count"][el])
                y[el, :n] = fp8.dequant_fp8(lref[d]["packed_recv_x"][el, :n], lref[d]["packed_recv_scales"][el, :n]) * (1 + el)
            lys.append(bits.f32_to_bf16(y.reshape(-1, H)).reshape(E // W, W * max_tok, H))
        pxh = u8(px)
        for el in range(E // W):
            n = int(lo["recv_count"][el])
            assert np.array_equal(pxh[el, :n], lo["packed_recv_x"][el, :n]), "ll fp8 bytes"
        lout, chook = buf.low_latency_combine(bits_to_bf16(lys[rank], dev), li[rank].to(dev), lw[rank].to(dev), lh,
                                              return_recv_hook=True)
        chook()
        lwant = combine.ll_combine(lys, [i.numpy() for i in li], [w.numpy() for w in lw], lref, E)[rank]
        assert np.array_equal(bf16_bits(lout), lwant), "ll combine"
    torch.cuda.synchronize()
    dist.barrier()
    print(f"GPU-MP-OK rank {rank}/{W}", flush=True)
    buf.destroy()
    dist.destroy_process_group()


if __name__ == "__main__":
    (run_cpu if sys.argv[1] == "cpu" else run_gpu)()
