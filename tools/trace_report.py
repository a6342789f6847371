"""Phase timelines from inside the kernels (B200EP_TRACE=1 makes them record %globaltimer stamps, read back with
b200ep_debug_trace).  Measurement aid, not part of the product:
    python -m torch.distributed.run --nproc-per-node N tools/trace_report.py [ll_tokens]
Stamp indices (ns since the first stamp of the kernel):
  LL dispatch   0 count CTA starts, 1 its counts are written | 2 token CTA 0 starts, 3 row quantised, 4 counts in and
                copies issued, 5 copies landed | last CTA: 6 all CTAs done, 7 system fence + flags out, 8 all sources' flags in
  LL combine    10 CTA 0 starts, 11 its rows are issued, 12 landed, 15 (the CTA that completes the last row) fence + flags out,
                13 all peers' flags seen, 14 reduced
  HT dispatch   20 CTA 0 starts, 21 layout kernel done (dependent launch), 22 all peers' counts in, 23 rows issued,
                24 copies landed, 25 signalled
  HT combine    30 CTA 0 starts, 31 chunk bounds ready, 40+c chunk c issued, 32 all landed and flagged,
                50+c all peers' flags of chunk c seen (warp 1), 60+c chunk c reduced (warp 1's share),
                70+c last driver arrived for chunk c, 80+c its system fence and flag stores are done
"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "llm-d_b200")):
    sys.path.insert(0, p)
os.environ["B200EP_TRACE"] = "1"
import builtins  # noqa: E402

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

_lines = []


def print(*a, **k):  # collected and written in one piece at the end: the ranks' lines do not interleave
    _lines.append(" ".join(str(x) for x in a))

import b200ep  # noqa: E402
from b200ep import LlHandle, _lib as L  # noqa: E402
from bench import E, H, K, T, make_inputs  # noqa: E402

LL_T = int(sys.argv[1]) if len(sys.argv) > 1 else 128
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
group = None
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
    group = dist.new_group(backend="gloo")
need = 2 * (world * T * (H + H // 32 + K * 12 + 4 + 352) + T * world * 2 * H) + (64 << 20)
buf = b200ep.Buffer(group, sym_bytes=int(max(need, b200ep.Buffer.ll_size_hint(LL_T, H, world, E)) * 1.1), device=local, rank=rank, world=world)


def barrier():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()


def stamps():
    a = (C.c_ulonglong * 128)()
    L.check(buf.lib.b200ep_debug_trace(buf.ctx, a), "debug_trace")
    return list(a)


def rel(s, idxs, base):
    return "  ".join(f"[{i}] {(s[i] - s[base]) / 1e3:7.1f}" for i in idxs if s[i])


x, ids, w = (t.to(dev) for t in make_inputs(rank))
ypool = torch.randn(world * T, H, device=dev).to(torch.bfloat16)
out = torch.empty(T, H, dtype=torch.bfloat16, device=dev)
for it in range(4):
    barrier()
    rx, rs, ri, rw, pe, h = buf.dispatch(x, ids, w, E, quantize=True)
    buf.combine(ypool[: h.num_recv], h, out)
barrier()
s = stamps()
if world > 1:
    print(f"rank {rank} HT dispatch us: {rel(s, range(20, 26), 20)}", flush=True)
    print(f"rank {rank} HT combine  us: {rel(s, [30, 31] + list(range(40, 48)) + [32], 30)}", flush=True)
    print(f"rank {rank} HT combine flags seen: {rel(s, range(50, 58), 30)}", flush=True)
    print(f"rank {rank} HT combine reduced:    {rel(s, range(60, 68), 30)}", flush=True)
    print(f"rank {rank} HT combine last arrival of chunk c (any CTA): {rel(s, range(70, 78), 30)}", flush=True)
    print(f"rank {rank} HT combine its fence + flags done:            {rel(s, range(80, 88), 30)}", flush=True)

g = torch.Generator().manual_seed(rank)
lx = torch.randn(LL_T, H, generator=g).to(torch.bfloat16).to(dev)
lids = torch.topk(torch.randn(LL_T, E, generator=g), K, dim=-1).indices.to(torch.int64).contiguous().to(dev)
lw = torch.rand(LL_T, K, generator=g).to(dev)
ly = torch.randn(E // world, world * LL_T, H, device=dev).to(torch.bfloat16)
lout = torch.empty(LL_T, H, dtype=torch.bfloat16, device=dev)
lh = LlHandle(0, LL_T, H, K, E, LL_T, True)
for it in range(4):
    barrier()
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    e0.record(); buf.ll_dispatch_launch(lh, lx, lids, False, L.LL_PHASE_ALL); e1.record()
    buf.low_latency_combine(ly, lids, lw, lh, lout, phases=L.PHASE_ALL); e2.record()
barrier()
s = stamps()
print(f"rank {rank} LL ({LL_T} tokens) events: dispatch {e0.elapsed_time(e1) * 1e3:.1f} us, combine {e1.elapsed_time(e2) * 1e3:.1f} us", flush=True)
print(f"rank {rank} LL dispatch us: {rel(s, range(0, 9), 0)}", flush=True)
print(f"rank {rank} LL combine  us: {rel(s, [10, 11, 12, 15, 13, 14], 10)}", flush=True)
barrier()
for r in range(world):
    if r == rank:
        builtins.print("\n".join(_lines), flush=True)
    barrier()
buf.destroy()
if world > 1:
    dist.destroy_process_group()
