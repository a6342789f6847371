#!/usr/bin/env python
"""Benchmark of the wide-EP MoE dispatch+combine hot path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One step = one high-throughput dispatch + combine of the synthetic DeepSeek-V3 MoE layer of
BASELINE.json configs[1] (hidden 7168, 256 experts, top-8, 4096 tokens/rank, bf16 -> FP8-128):
routing layout kernel, count exchange, fused quantise+dispatch, combine.  EP size = N (one
process per GPU, launched by torchrun for N > 1; CUDA-IPC peers, NCCL only for bootstrap/barriers).

Printed JSON (one line, rank 0):
  value      whole-job tokens/s, inputs resident in HBM, CUDA-event timed, max over ranks
  e2e        same metric through the public API with HOST buffers: pinned->device copies of the
             step's inputs and a device->host read of the combined output inside the timed region
             (double-buffered on copy streams; expert outputs are resident bf16 rows)
  roofline   dominant kernel vs measured HBM peak (N=1) or NVLink per-direction peak (N>1)
  cpu_baseline  the reference's CPU dispatch/combine (oracle/cpu_reference.py, a port of vLLM
             cpu_fused_moe_torch) timed on this box's host cores on a bounded sample (N=1 only)
  ll         low-latency decode mode, 128 tokens/rank, CUDA-graph replay p50/p99 (informative)
--impl reference times that CPU port alone, all host threads, and prints the same line shape.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "llm-d_b200"), os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402

T, H, E, K = 4096, 7168, 256, 8
LL_T = 128
NVLINK_PEAK_GBS = 770.0  # measured peer-copy GB/s per direction on this pool (B200_PROFILING.md); nominal 900
METRIC = "dispatch+combine tokens/s"
DTYPE = "bf16 in, fp8 e4m3 + fp32 scales dispatched, bf16 combined with fp32 accumulate"
WORKLOAD = "synthetic DeepSeek-V3 MoE layer (hidden 7168, 256 experts, top-8), 4096 tokens/rank bf16->FP8 high-throughput dispatch+combine"


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "MEASURED_PEAKS.json"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback (B200_PROFILING.md)"


def make_inputs(seed: int, tokens: int = T):
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(tokens, H, generator=g) * (torch.rand(tokens, 1, generator=g) * 4)).to(torch.bfloat16)
    logits = torch.randn(tokens, E, generator=g)
    vals, ids = torch.topk(logits, K, dim=-1, sorted=False)
    w = torch.softmax(vals, dim=-1).float()
    return x.contiguous(), ids.to(torch.int64).contiguous(), w.contiguous()


_SAMPLER_SRC = r"""
import sys, time
import pynvml as nv
nv.nvmlInit()
h = nv.nvmlDeviceGetHandleByIndex(int(sys.argv[1]))
period = float(sys.argv[2])
print("max", nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM), flush=True)
get_r = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
while True:
    print(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM), int(get_r(h)), flush=True)
    time.sleep(period)
"""


class ClockSampler:
    """SM clock and throttle reasons of rank 0's GPU, sampled DURING the timed region by a helper
    PROCESS that polls NVML (two light queries every 10 ms).

    Tried first and rejected because they slowed rank 0 down, and with it every rank that waits
    for it in the count exchange: a looping `nvidia-smi -lms 100` (~1 ms per step at N=4), and
    NVML polled from a thread of the benchmark process itself (initialising NVML in-process made
    each kernel launch ~20 us slower: +150 us per step at N=2)."""

    def __init__(self, index: int, period_s: float = 0.01):
        self.index, self.period, self.lines, self.proc = index, period_s, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen([sys.executable, "-c", _SAMPLER_SRC, str(self.index), str(self.period)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            first = self.proc.stdout.readline()  # wait until NVML is up, outside the timed region
            self.mx = float(first.split()[1]) if first.startswith("max") else None
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append((time.perf_counter(), line))

    def mark(self):
        return time.perf_counter()

    def stop(self, t0=None, t1=None):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml helper unavailable"]}
        time.sleep(2.5 * self.period)
        self.proc.terminate()
        sm, bits = [], 0
        for ts, ln in self.lines:
            if t0 is not None and not (t0 <= ts <= t1 + 2 * self.period):
                continue
            p = ln.split()
            if len(p) == 2:
                sm.append(float(p[0])); bits |= int(p[1])
        sm.sort()
        names = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
                 0x80: "hw_power_brake_slowdown"}
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.mx,
                "reasons": sorted(n for b, n in names.items() if bits & b), "samples": len(sm),
                "how": "NVML polled every 10 ms by a helper process while the timed steps ran"}


def cpu_baseline(sample_tokens: int, repeats: int, threads: int | None = None):
    """The reference's CPU dispatch/combine on a bounded sample of the same workload."""
    from oracle.cpu_reference import cpu_dispatch_combine

    n = threads or os.cpu_count() or 1
    torch.set_num_threads(n)
    x, ids, w = make_inputs(1234, sample_tokens)
    ids32 = ids.to(torch.int32)
    cpu_quant_fp8(x); cpu_dispatch_combine(x, w, ids32, E)  # warm-up
    best, times = None, []
    for _ in range(repeats):
        t0 = time.perf_counter()
        cpu_quant_fp8(x)
        cpu_dispatch_combine(x, w, ids32, E)
        dt = time.perf_counter() - t0
        times.append(dt)
        best = dt if best is None else min(best, dt)
    return {
        "value": sample_tokens / best,
        "unit": "tokens/s",
        "cores": torch.get_num_threads(),
        "kind": "port",
        "sample": f"{sample_tokens} tokens x hidden {H}, top-{K} of {E} experts, bf16 -> FP8-128 quantise, identity experts, "
                  f"best of {repeats} (mean {sum(times) / len(times):.3f} s/step); "
                  "port of vLLM cpu_fused_moe_torch (cpu_fused_moe.py:444-449,467-481)",
    }, best


NBUF = 4  # rotating input sets, so that a timed step never finds its inputs in the 126 MB L2


def _config(world: int):
    """The `config` object: identical in the b200ep arm and in the reference arm (the driver compares them)."""
    return {"workload": WORKLOAD, "ep_size": world, "tokens_per_rank": T, "hidden": H, "experts": E, "topk": K,
            "l2": f"GPU arm: {NBUF} rotating input sets ({NBUF * (T * H * 2) >> 20} MiB of x) > 126 MB L2; host arm: not applicable"}


def _profile_json(name):
    """Numbers that come from committed profiler output (profiles/), never literals in this file."""
    try:
        with open(os.path.join(ROOT, "profiles", name)) as f:
            return json.load(f)
    except Exception:
        return None


def cpu_quant_fp8(x: torch.Tensor):
    """The reference's CPU form of per_token_group_quant_fp8 (vLLM fp8_utils.py:510-591, group 128): what its
    high-throughput path runs before the dispatch (deepep_ht.py:282-289), in torch on the host cores."""
    g = x.float().view(x.shape[0], -1, 128)
    s = g.abs().amax(dim=2).clamp_(min=1e-10) / 448.0
    q = (g / s.unsqueeze(-1)).clamp_(-448.0, 448.0).to(torch.float8_e4m3fn)
    return q.view(x.shape[0], -1), s


def run_reference(args):
    """The reference's own CPU implementation of the path (oracle/cpu_reference.py, a port of vLLM's
    cpu_fused_moe_torch: the reference repo holds no source for it), all host threads.  One step = the SAME
    workload as a b200ep step at this N: N ranks' worth of 4096 tokens (the CPU has no ranks: the expert-parallel
    exchange degenerates to the gather/scatter over all N*4096 tokens), bf16 in, FP8-128 quantise, dispatch,
    identity experts, fp32 weighted combine."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    steps, warm = max(1, args.steps), max(0, args.warmup)
    from oracle.cpu_reference import cpu_dispatch_combine

    world = max(1, args.gpus)
    torch.set_num_threads(os.cpu_count() or 1)
    parts = [make_inputs(1234 + r) for r in range(world)]
    x = torch.cat([p[0] for p in parts]); ids32 = torch.cat([p[1] for p in parts]).to(torch.int32)
    w = torch.cat([p[2] for p in parts])
    tokens = x.shape[0]

    def step(stable=False):
        nonlocal x, ids32, w
        cpu_quant_fp8(x)  # the payload the reference dispatches is FP8; the CPU fragment itself moves bf16 rows
        return cpu_dispatch_combine(x, w, ids32, E, stable=stable)

    # bounded: if K steps of the full workload would take more than ~2.5 minutes on these cores, a step becomes a
    # proportionally smaller sample of it (tokens/s is what is reported either way)
    p0 = time.perf_counter(); step(); probe = time.perf_counter() - p0
    if probe * (steps + warm) > 150.0:
        keep = max(256, int(tokens * 150.0 / (probe * (steps + warm))) // 256 * 256)
        x, ids32, w = x[:keep].contiguous(), ids32[:keep].contiguous(), w[:keep].contiguous()
        tokens = keep
    for _ in range(warm):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    t1 = time.perf_counter(); step(stable=True); dt_stable = time.perf_counter() - t1
    t2 = time.perf_counter(); cpu_quant_fp8(x); dt_quant = time.perf_counter() - t2
    val = tokens / dt
    base = {"value": val, "unit": "tokens/s", "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"{tokens} of {world * T} tokens/step x {steps} steps, hidden {H}, top-{K} of {E}, bf16 -> FP8-128 quantise + dispatch + "
                      f"identity experts + fp32 combine; of which quantise {dt_quant * 1e3:.0f} ms; the stable-argsort variant "
                      f"takes {dt_stable * 1e3:.0f} ms/step"}
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": "tokens/s", "n_gpus": args.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": DTYPE, "data": "synthetic", "config": _config(world),
        "cpu_baseline": base, "e2e": {"value": val, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))
    return 0


def run_nccl_agrs(args):
    """On-box GPU comparator: the dense NCCL stand-in the reference CI substitutes for
    DeepEP -- AgRsAll2AllManager (vllm/distributed/device_communicators/all2all.py:41-140): dispatch =
    all-gather of every rank's fp8 tokens + scales + top-k ids/weights, combine = reduce-scatter of the
    bf16 [W*T, H] partials.  Collectives only (no quantise, no experts), same shapes, CUDA events."""
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    steps, warm = max(1, args.steps), max(3, args.warmup)
    xq = torch.randint(0, 255, (T, H), dtype=torch.uint8, device=dev)
    xs = torch.randn(T, H // 128, device=dev)
    ids = torch.randint(0, E, (T, K), device=dev)
    w = torch.rand(T, K, device=dev)
    g = [torch.empty((world * t.shape[0],) + tuple(t.shape[1:]), dtype=t.dtype, device=dev) for t in (xq, xs, ids, w)]
    y = torch.randn(world * T, H, device=dev).to(torch.bfloat16)
    out = torch.empty(T, H, dtype=torch.bfloat16, device=dev)

    def step():
        if world > 1:
            for src, dst in zip((xq, xs, ids, w), g):
                dist.all_gather_into_tensor(dst, src)
            dist.reduce_scatter_tensor(out, y)
        else:
            g[0].copy_(xq); out.copy_(y[:T])

    for _ in range(warm):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        t = float(ms.item()) / steps
        print(json.dumps({"impl": "nccl_agrs", "metric": METRIC, "value": world * T / (t * 1e-3), "unit": "tokens/s",
                          "n_gpus": world, "steps": steps, "warmup": warm, "ms_per_step": t, "higher_is_better": True,
                          "scaling": "weak", "data": "synthetic",
                          "config": {"workload": WORKLOAD, "ep_size": world,
                                     "what": "NCCL all_gather (fp8+scales+ids+weights) + reduce_scatter (bf16), collectives only",
                                     "tokens_per_s_per_gpu": T / (t * 1e-3)}}))
    if world > 1:
        dist.destroy_process_group()
    return 0


def run_flashinfer_a2a(args):
    """On-box GPU comparator: FlashInfer's one-sided NVLink MoE all-to-all (flashinfer.comm.MoeAlltoAll,
    the kernels behind vLLM's ``--all2all-backend flashinfer_nvlink_one_sided``), LIBRARY code, called the way vLLM calls
    it (prepare_finalize/flashinfer_nvlink_one_sided.py:119-168): dispatch of fp8 rows + fp32 scales + int32 ids + fp32
    weights, combine of bf16 rows; the workspace is FlashInfer's own (cuMem + POSIX fd exchange over the gloo group).
    Same shapes as the b200ep arm: T=4096 per rank (throughput) and T=32..256 (decode, CUDA-graph replay when the ops
    capture, else back-to-back launches).  Its JIT module is prebuilt for sm_100a under baseline/_fi_cache (no network)."""
    os.environ.setdefault("FLASHINFER_WORKSPACE_BASE", os.path.join(ROOT, "baseline", "_fi_cache"))
    os.environ.setdefault("FLASHINFER_CUDA_ARCH_LIST", "10.0a")
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    def emit(obj):
        if rank == 0:
            os.write(real_stdout, (json.dumps(obj) + "\n").encode())

    if world < 2:
        emit({"impl": "flashinfer_a2a", "unavailable": "needs at least 2 GPUs (torchrun --nproc-per-node N)"})
        return 0
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    cpu_group = dist.new_group(backend="gloo")
    try:
        from flashinfer.comm import MoeAlltoAll
        from flashinfer.comm.mapping import Mapping
        from flashinfer.comm.mnnvl import MnnvlConfig, TorchDistBackend
        from flashinfer.comm.trtllm_moe_alltoall import moe_a2a_get_workspace_size_per_rank

        max_tok = T
        disp_bytes = H + 4 * (H // 128) + K * 4 + K * 4
        ws = moe_a2a_get_workspace_size_per_rank(ep_size=world, max_num_tokens=max_tok,
                                                 total_dispatch_payload_size_per_token=disp_bytes,
                                                 combine_payload_size_per_token=2 * H)
        mapping = Mapping(world, rank, torch.cuda.device_count(), tp_size=world, moe_ep_size=world)
        a2a = MoeAlltoAll(mapping=mapping, max_num_tokens=max_tok, top_k=K, num_experts=E, workspace_size_per_rank=ws,
                          mnnvl_config=MnnvlConfig(comm_backend=TorchDistBackend(cpu_group)))
    except Exception as ex:
        emit({"impl": "flashinfer_a2a", "unavailable": f"FlashInfer MoeAlltoAll could not be set up on this box: {ex!r}"[:400]})
        dist.destroy_process_group()
        return 0
    dist.barrier()

    def make(tokens):
        g = torch.Generator(device="cpu").manual_seed(1000 + rank)
        xq = torch.randint(0, 255, (tokens, H), dtype=torch.uint8, generator=g).to(dev).view(torch.float8_e4m3fn)
        xs = torch.rand(tokens, H // 128, generator=g).to(dev)
        _, ids = torch.topk(torch.randn(tokens, E, generator=g), K, dim=-1)
        w = torch.rand(tokens, K, generator=g).to(dev)
        return xq, xs, ids.to(torch.int32).to(dev), w

    def one(tokens, inp, y):
        xq, xs, ids, w = inp
        a2a.dispatch(token_selected_experts=ids, input_payloads=[xq, xs, ids, w], runtime_max_tokens_per_rank=tokens,
                     invalid_token_expert_id=-1, expert_id_payload_index=2)
        return a2a.combine(payload=y, runtime_max_tokens_per_rank=tokens)

    def timed(tokens, iters, warm, graph):
        inp = make(tokens)
        y = torch.randn(world, tokens, H, device=dev).to(torch.bfloat16)
        for _ in range(warm):
            one(tokens, inp, y)
        torch.cuda.synchronize(); dist.barrier()
        g, how = None, "launches queued back to back"
        if graph:
            try:
                st = torch.cuda.Stream(device=dev)
                with torch.cuda.stream(st):
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g, stream=st, capture_error_mode="relaxed"):
                        one(tokens, inp, y)
                how = "CUDA graph of dispatch + combine, replayed"
            except Exception:
                g = None
                a2a._state.phase = "idle"
            torch.cuda.synchronize()
        flag = torch.tensor([1 if g is not None else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if not int(flag.item()):
            g, how = None, "launches queued back to back"
        dist.barrier()
        samples = []
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
        for a, b in evs:
            a.record()
            g.replay() if g is not None else one(tokens, inp, y)
            b.record()
        torch.cuda.synchronize()
        samples = sorted(a.elapsed_time(b) for a, b in evs)
        t = torch.tensor([samples[len(samples) // 2], sum(samples) / len(samples)], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0].item()), float(t[1].item()), how

    steps, warm = max(1, args.steps), max(3, args.warmup)
    try:
        _, ht_mean, _ = timed(T, steps, warm, graph=False)
        ll = {}
        for tok in (32, 64, 128, 256):
            p50, _, how = timed(tok, 200, 10, graph=True)
            ll[str(tok)] = {"p50_us": p50 * 1e3, "how": how}
    except Exception as ex:
        emit({"impl": "flashinfer_a2a", "unavailable": f"FlashInfer MoeAlltoAll failed while running: {ex!r}"[:400]})
        os._exit(0)
    emit({"impl": "flashinfer_a2a", "metric": METRIC, "value": world * T / (ht_mean * 1e-3), "unit": "tokens/s", "n_gpus": world,
          "steps": steps, "warmup": warm, "ms_per_step": ht_mean, "higher_is_better": True, "scaling": "weak", "data": "synthetic",
          "config": {"workload": WORKLOAD, "ep_size": world, "tokens_per_s_per_gpu": T / (ht_mean * 1e-3),
                     "what": "flashinfer.comm.MoeAlltoAll dispatch (fp8 rows + fp32 scales + int32 ids + fp32 weights, already "
                             "quantised: no quantise pass, no layout kernel) + combine (bf16), called as vLLM's "
                             "flashinfer_nvlink_one_sided prepare/finalize does; library kernels, comparator only"},
          "ll": ll, "deep_ep": "unavailable on this box (import deep_ep finds only the b200ep shim; DeepEP needs NVSHMEM)"})
    torch.cuda.synchronize()
    os._exit(0)  # FlashInfer's workspace teardown can hang at interpreter exit; the numbers are out


def run_probe(args):
    """What NVLink gives SM-driven copies on this box: every rank
    moves 32 MiB to / from each peer between the symmetric regions, in pieces of the sizes the kernels use
    (one fp8 row 7 KB, one bf16 row 14 KB, several rows), by TMA bulk copies, by register copies, by peer loads,
    and with the copy engines.  GB/s per direction per GPU, CUDA events, max time over ranks.  Needs N >= 2."""
    import ctypes as C

    import torch.distributed as dist

    import b200ep
    from b200ep import _lib as L

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world < 2:
        print(json.dumps({"impl": "probe", "unavailable": "needs at least 2 GPUs (torchrun --nproc-per-node N)"}))
        return 0
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    group = dist.new_group(backend="gloo")
    bpp = 32 << 20
    buf = b200ep.Buffer(group, sym_bytes=2 * world * bpp + (4 << 20), device=local, rank=rank, world=world)
    probe = C.CDLL(os.path.join(os.path.dirname(L.LIB_PATH), "libb200ep_probe.so"))
    probe.b200ep_probe_run.restype = C.c_int
    probe.b200ep_probe_run.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int, C.c_void_p]
    bases = (C.c_void_p * world)()
    for r in range(world):
        p = C.c_void_p()
        L.check(buf.lib.b200ep_peer_base(buf.ctx, r, C.byref(p)), "peer_base")
        bases[r] = p.value + (1 << 20)  # skip the control block
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    names = {0: "push_tma", 1: "push_regs", 2: "pull_regs", 3: "pull_tma", 4: "push_copy_engine",
             5: "push_tma_random_dst", 6: "push_tma_fanout", 7: "push_tma_fanout_small"}
    cases = []
    for mode in (0, 3):
        for piece in (7168, 14336, 28672, 43008):
            cases.append((mode, piece, sms))
        cases.append((mode, 14336, 2 * sms))
        cases.append((mode, 7168, 3 * sms))
    for mode in (1, 2):
        for piece in (7168, 14336, 65536):
            cases.append((mode, piece, 2 * sms))
    cases.append((4, 16, 1))
    for mode in (5, 6, 7):
        cases.append((mode, 7168, sms))
        cases.append((mode, 7168, 3 * sms))
    if args.probe_modes:
        want = {int(m) for m in args.probe_modes.split(",")}
        cases = [c for c in cases if c[0] in want]
    st = torch.cuda.current_stream().cuda_stream
    R = 4
    out = []
    for mode, piece, grid in cases:
        def run():
            rc = probe.b200ep_probe_run(bases, world, rank, mode, bpp, piece, grid, st)
            assert rc == 0, f"probe mode {mode} failed: {rc}"
        run(); run()
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(R):
            run()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / R], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.barrier()
        moved = (world - 1) * bpp
        if mode in (6, 7):  # pieces * expected copies (2/3 of the peers), see probe_fanout_kernel
            npieces = (bpp // piece) * (world - 1) * 3 // (2 * (world - 1)) + 1
            moved = npieces * piece * (world - 1) * 2 / 3
        gbs = moved / (float(t.item()) * 1e-3) / 1e9
        out.append({"mode": names[mode], "piece_bytes": piece if mode != 4 else bpp, "ctas": grid, "ms": float(t.item()), "gbs_per_direction": gbs})
        if rank == 0:
            print(f"[probe] {names[mode]:17s} piece {piece:6d} ctas {grid:4d}: {gbs:7.1f} GB/s", file=sys.stderr, flush=True)
    if rank == 0:
        print(json.dumps({"impl": "probe", "n_gpus": world, "bytes_per_peer": bpp, "unit": "GB/s per direction per GPU",
                          "how": "every rank moves 32 MiB to/from each peer at once (permutation order), CUDA events, max over ranks",
                          "results": out}))
    torch.cuda.synchronize(); dist.barrier()
    buf.destroy()
    dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200ep")
    ap.add_argument("--no-ll", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--sym-gb", type=float, default=0.0)
    ap.add_argument("--breakdown", action="store_true", help="per-phase CUDA-event breakdown of the step (stderr)")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle check of the timed calls (experiments only)")
    ap.add_argument("--probe-modes", default="", help="--impl probe: comma-separated mode numbers to run (default all)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if args.impl == "nccl_agrs":
        return run_nccl_agrs(args)
    if args.impl == "probe":
        return run_probe(args)
    if args.impl == "flashinfer_a2a":
        return run_flashinfer_a2a(args)

    import torch.distributed as dist

    import b200ep
    from b200ep import _lib as L

    # libraries (NCCL's version banner) write to stdout; the contract is ONE JSON line there: everything else goes to stderr
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world} (launch with torchrun for N>1)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        group = dist.new_group(backend="gloo")  # host channel for the 64-byte IPC handles

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    steps, warm = max(1, args.steps), max(3, args.warmup)
    W = world
    # per-slot need: W*T receive rows + T*W stage rows; size the region for it with headroom
    need = 2 * (W * T * (H + H // 32 + K * 12 + 4) + T * W * 2 * H) + (64 << 20)
    ll_need = max(b200ep.Buffer.ll_size_hint(256, H, W, E), b200ep.Buffer.ll_size_hint(256, H, W, 288) if 288 % W == 0 else 0)
    sym = int(args.sym_gb * 2**30) if args.sym_gb else int(max(need, ll_need) * 1.1)
    buf = b200ep.Buffer(group, sym_bytes=sym, device=local, rank=rank, world=world)

    # ---- parity first: one untimed step through exactly the calls timed below (Buffer.dispatch + Buffer.combine with
    # all phases, the host-sync-free pair, graph-replayed low-latency dispatch + combine), every rank's outputs compared
    # bit for bit with the CPU oracle.  A mismatch aborts the run on every rank: no number without parity.
    parity_res = None
    if not args.no_parity:
        import parity

        try:
            parity_res = parity.run_bench_parity(buf, group, world, rank, dev, barrier, H=H, E=E, K=K, T_ht=1100, T_ll=LL_T,
                                                 ll_experts=(E, 288))
        except parity.ParityError as ex:
            print(f"[rank {rank}] PARITY FAILED: {ex}", file=sys.stderr, flush=True)
            if rank == 0:
                os.write(real_stdout, (json.dumps({"metric": METRIC, "value": None, "n_gpus": world, "parity": {"error": str(ex)[:2000]}}) + "\n").encode())
            os._exit(3)
        barrier()

    xs, idss, wss = [], [], []
    for b in range(NBUF):
        x, ids, w = make_inputs(1000 * b + rank)
        xs.append(x.to(dev)); idss.append(ids.to(dev)); wss.append(w.to(dev))
    y_pool = torch.randn(W * T if W > 1 else T, H, device=dev).to(torch.bfloat16)  # expert outputs (resident)
    out = torch.empty(T, H, dtype=torch.bfloat16, device=dev)

    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    remote_copies = []
    last = {}

    marks = []

    def step(i, timed):
        b = i % NBUF
        if args.breakdown and timed:
            m = [ev() for _ in range(3)]
            th0 = time.perf_counter()
            m[0].record()
        # the public API: layout + count exchange (host wait) + fused quantise/dispatch in one call
        rx, rs, ridx, rw, per_expert, h = buf.dispatch(xs[b], idss[b], wss[b], E, quantize=True)
        if args.breakdown and timed:
            th1 = time.perf_counter()
            m[1].record()
        y = y_pool[: h.num_recv]
        buf.combine(y, h, out)
        if args.breakdown and timed:
            m[2].record(); marks.append((m, th1 - th0, time.perf_counter() - th1))
        last["h"], last["a"], last["y"], last["pe"] = h, (xs[b], None, idss[b], wss[b], False), y, per_expert
        if timed:
            remote_copies.append(h.num_recv - (h.rank_prefix[rank + 1] - h.rank_prefix[rank]))

    for i in range(warm):
        step(i, False)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    t_start, t_end = ev(), ev()
    t_start.record()
    th0 = time.perf_counter()
    for i in range(steps):
        step(warm + i, True)
    t_end.record()
    torch.cuda.synchronize()
    th1 = time.perf_counter()
    ms = t_start.elapsed_time(t_end)
    clocks = sampler.stop(th0, th1) if rank == 0 else None
    tms = torch.tensor([ms], device=dev)
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    barrier()
    ms_total = float(tms.item())
    ms_step = ms_total / steps
    value = world * T * steps / (ms_total / 1e3)

    # ---- the same step without the host round trip (num_worst_tokens mode): informative
    worst = (W * T) if W > 1 else T
    yw = y_pool[:worst]

    def step_nosync(i):
        b = i % NBUF
        _, _, _, _, _, hh = buf.dispatch(xs[b], idss[b], wss[b], E, quantize=True, num_worst_tokens=worst)
        buf.combine(yw, hh, out)

    for i in range(3):
        step_nosync(i)
    barrier()
    a0, a1 = ev(), ev()
    a0.record()
    for i in range(steps):
        step_nosync(i)
    a1.record()
    torch.cuda.synchronize()
    ams = torch.tensor([a0.elapsed_time(a1)], device=dev)
    if world > 1:
        dist.all_reduce(ams, op=dist.ReduceOp.MAX)
    barrier()
    nosync = {"value": world * T * steps / (float(ams.item()) / 1e3), "unit": "tokens/s",
              "ms_per_step": float(ams.item()) / steps,
              "what": "dispatch(num_worst_tokens=W*T) + combine: no host wait, per-expert counts stay on the device"}
    # restore a synchronous handle for the per-kernel section below
    step(0, False)
    torch.cuda.synchronize()

    # ---- per-kernel durations: the same launches, back to back (no host round trip in between),
    # one CUDA-event pair per launch, L2 flushed before each by READING a 512 MB buffer on the same
    # stream (a memset would leave L2 full of dirty lines whose write-back is then billed to the kernel)
    flush = torch.zeros(512 << 20, dtype=torch.uint8, device=dev)
    R = 10
    k_disp, k_comb = [], []
    h, a, y = last["h"], last["a"], last["y"]
    barrier()
    k_dsend, k_csend = [], []
    for _ in range(R):
        flush.amax()
        e0, e1, e2 = ev(), ev(), ev()
        e0.record(); buf.ht_dispatch_launch(h, *a, L.PHASE_SEND); e1.record()
        buf.ht_dispatch_launch(h, *a, L.PHASE_RECV); e2.record()
        k_disp.append((e0, e2)); k_dsend.append((e0, e1))
    for _ in range(R):
        flush.amax()
        e0, e1, e2 = ev(), ev(), ev()
        e0.record(); buf.combine(y, h, out, L.PHASE_SEND); e1.record()
        buf.combine(y, h, out, L.PHASE_RECV); e2.record()
        k_comb.append((e0, e2)); k_csend.append((e0, e1))
    torch.cuda.synchronize()
    barrier()
    # ---- expert-grouped layout: regroup the received rows by local expert (replaces vLLM's ep_scatter)
    # and combine straight from the grouped GEMM output with the weighted fp32 reduction inside the push (replaces
    # ep_gather + Buffer.combine).  Informative: event-timed like the kernels above, max over ranks.
    grouped = None
    try:
        if parity_res is not None and parity_res.get("ht_grouped") != "ok":
            raise RuntimeError("the grouped-layout calls failed their oracle check (see parity.ht_grouped): not timed")
        a_perm, s_perm, eids, inv = buf.ht_permute(h, per_expert=last["pe"])
        y_grp = torch.randn(a_perm.shape[0], H, device=dev).to(torch.bfloat16)
        g_perm, g_comb = [], []
        for _ in range(R):
            flush.amax()
            e0, e1 = ev(), ev()
            e0.record(); buf.ht_permute(h, per_expert=last["pe"], out_x=a_perm, out_scales=s_perm); e1.record()
            g_perm.append((e0, e1))
        for _ in range(R):
            flush.amax()
            e0, e1 = ev(), ev()
            e0.record(); buf.combine_permuted(y_grp, inv, h, out, phases=L.PHASE_SEND)
            buf.combine_permuted(y_grp, inv, h, out, phases=L.PHASE_RECV); e1.record()
            g_comb.append((e0, e1))
        torch.cuda.synchronize()
        gt = torch.tensor([sum(p.elapsed_time(q) for p, q in g_perm) / R, sum(p.elapsed_time(q) for p, q in g_comb) / R],
                          device=dev)
        if world > 1:
            dist.all_reduce(gt, op=dist.ReduceOp.MAX)
        entries = int(sum(last["pe"]))
        perm_bytes = (h.num_recv + entries) * (H + 4 * H // 128)
        grouped = {
            "permute_ms": float(gt[0].item()), "rows_in": h.num_recv, "rows_out": entries, "m_sum": int(a_perm.shape[0]),
            "permute_algorithmic_bytes": perm_bytes, "permute_hbm_gbs": perm_bytes / (float(gt[0].item()) * 1e-3) / 1e9,
            "combine_permuted_ms": float(gt[1].item()),
            "what": "rank 0's row counts; permute = plan + scatter kernels (rows read once, written once per local expert); "
                    "combine_permuted = gather + fp32 weighted reduce + push, then the fp32 reduce on the source (two launches)",
        }
        del y_grp, a_perm, s_perm
    except Exception as ex:  # informative block: never lose the headline over it
        grouped = {"error": repr(ex)[:300]}
    barrier()
    del flush
    if args.breakdown:
        n = len(marks)
        seg = [sum(m[j].elapsed_time(m[j + 1]) for m, _, _ in marks) / n * 1e3 for j in range(2)]
        print(f"[rank {rank}] step breakdown us: layout+notify+dispatch (gpu) {seg[0]:.1f} | combine (gpu) {seg[1]:.1f} || "
              f"host in dispatch() {sum(a for _, a, _ in marks) / n * 1e6:.1f} host in combine() {sum(b for _, _, b in marks) / n * 1e6:.1f}",
              file=sys.stderr, flush=True)
    d_ms = sorted(p.elapsed_time(q) for p, q in k_disp)
    c_ms = sorted(p.elapsed_time(q) for p, q in k_comb)
    d_avg, c_avg = sum(d_ms) / len(d_ms), sum(c_ms) / len(c_ms)
    kt = torch.tensor([d_avg, c_avg], device=dev)
    if world > 1:
        dist.all_reduce(kt, op=dist.ReduceOp.MAX)
    d_avg, c_avg = float(kt[0].item()), float(kt[1].item())
    ds_avg = sum(p.elapsed_time(q) for p, q in k_dsend) / R
    cs_avg = sum(p.elapsed_time(q) for p, q in k_csend) / R
    peaks, peak_src = measured_peaks()
    if world == 1:
        # algorithmic bytes per launch: dispatch T*(2H + H + 4H/128); combine T*(2H + 2H)
        db, cb = T * (2 * H + H + 4 * H // 128), T * (2 * H + 2 * H)
        peak, bound, unit_note = float(peaks["hbm_gbs"]), "hbm", f"HBM copy peak from {peak_src}"
    else:
        r = sum(remote_copies) / len(remote_copies)  # remote copies actually made per step by this rank's peers -> me
        db, cb = r * (H + 4 * H // 128), r * 2 * H
        probe = _profile_json("r02_fabric_probe.json") or {}
        if str(world) in probe.get("best_push_gbs", {}):
            peak = float(probe["best_push_gbs"][str(world)])
            unit_note = (f"best SM-driven push between the symmetric regions measured on this pool at EP={world} with every rank "
                         "pushing at once (bench.py --impl probe, profiles/r02_fabric_probe.json); 900 nominal, 770 one-directional peer copy")
        else:
            peak, unit_note = NVLINK_PEAK_GBS, "measured peer-copy GB/s per direction (guide); 900 nominal"
        bound = "nvlink"
    # roofline durations: the kernels that move the algorithmic bytes.  N>1: the push kernels (their
    # duration does not include waiting for peers); N=1: dispatch send kernel and the reduce kernel.
    st = torch.tensor([ds_avg, cs_avg], device=dev)
    if world > 1:
        dist.all_reduce(st, op=dist.ReduceOp.MAX)
    ds_max, cs_max = float(st[0].item()), float(st[1].item())
    d_k, c_k = (ds_max, cs_max) if world > 1 else (ds_avg, c_avg)
    d_gbs, c_gbs = db / (d_k * 1e-3) / 1e9, cb / (c_k * 1e-3) / 1e9
    if world > 1:
        dominant = "ht_combine_send_kernel (bf16 rows, TMA load -> TMA store to the peer)" if c_k >= d_k else "ht_dispatch_send_kernel (fused quantise + TMA fan-out)"
    else:
        dominant = "ht_combine_local_kernel (gather, fp32 accumulate)" if c_k >= d_k else "ht_dispatch_local_kernel (fused quantise)"
    dom_gbs = c_gbs if c_k >= d_k else d_gbs
    ncu = _profile_json("r02_ncu_n1.json") if world == 1 else None
    traffic = None
    if ncu:
        traffic = ncu.get("combine_dram_bytes") if c_k >= d_k else ncu.get("dispatch_dram_bytes")
    roofline = {
        "bound": bound, "achieved": dom_gbs, "peak": peak, "unit": "GB/s", "frac": dom_gbs / peak,
        # dram__bytes_read + dram__bytes_write per launch of that kernel, read from the committed summary of the
        # ncu --set full capture (profiles/r02_ncu_n1.json, made by tools/ncu_summary.py); null when there is none
        "traffic": traffic,
        "kernel": dominant, "peak_source": unit_note,
        "how": "CUDA events around each kernel, launches queued back to back, L2 flushed (512 MB read) before each; "
               "N>1: push kernels only (max over ranks), N=1: dispatch send kernel and reduce kernel",
        "kernels": {
            "dispatch": {"ms": d_k, "send_plus_wait_ms": d_avg, "send_kernel_ms": ds_avg, "algorithmic_bytes": db,
                         "gbs": d_gbs, "frac": d_gbs / peak},
            "combine": {"ms": c_k, "send_plus_reduce_ms": c_avg, "send_kernel_ms": cs_avg, "algorithmic_bytes": cb,
                        "gbs": c_gbs, "frac": c_gbs / peak},
        },
        "step_frac": (db + cb) / (ms_step * 1e-3) / 1e9 / peak,
    }

    # ---- end to end through the public API with host buffers: every step copies ITS inputs from pinned host memory
    # and ITS combined output back, all inside the timed region; the copies run on streams of their own with two
    # buffer sets, so step i+1's inputs travel while step i computes and step i-1's output returns (PCIe is full
    # duplex).  Expert outputs are resident bf16 rows, as in the device-timed step (the GEMMs are out of scope).
    hx = [x.cpu().pin_memory() for x in xs[:2]]
    hi = [i.cpu().pin_memory() for i in idss[:2]]
    hw = [w.cpu().pin_memory() for w in wss[:2]]
    houts = [torch.empty(T, H, dtype=torch.bfloat16).pin_memory() for _ in range(2)]
    dxs = [torch.empty_like(xs[0]) for _ in range(2)]
    dis = [torch.empty_like(idss[0]) for _ in range(2)]
    dws = [torch.empty_like(wss[0]) for _ in range(2)]
    douts = [torch.empty(T, H, dtype=torch.bfloat16, device=dev) for _ in range(2)]
    s_in, s_out, s_main = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev), torch.cuda.current_stream()
    ev_in = [torch.cuda.Event() for _ in range(2)]
    ev_comp = [torch.cuda.Event() for _ in range(2)]
    ev_out = [torch.cuda.Event() for _ in range(2)]
    for e in ev_comp + ev_out:
        e.record(s_main)

    def e2e_h2d(i):
        b = i % 2
        with torch.cuda.stream(s_in):
            s_in.wait_event(ev_comp[b])  # the step that last used this input set has finished with it
            dxs[b].copy_(hx[b], non_blocking=True); dis[b].copy_(hi[b], non_blocking=True); dws[b].copy_(hw[b], non_blocking=True)
            ev_in[b].record(s_in)

    def e2e_step(i):
        b = i % 2
        s_main.wait_event(ev_in[b])
        s_main.wait_event(ev_out[b])  # this output buffer's previous copy to the host is done
        rx, rs, ridx, rw, per_expert, h = buf.dispatch(dxs[b], dis[b], dws[b], E, quantize=True)
        buf.combine(y_pool[: h.num_recv], h, douts[b])
        ev_comp[b].record(s_main)
        with torch.cuda.stream(s_out):
            s_out.wait_event(ev_comp[b])
            houts[b].copy_(douts[b], non_blocking=True)
            ev_out[b].record(s_out)

    def e2e_run(n):
        e2e_h2d(0)
        for i in range(n):
            if i + 1 < n:
                e2e_h2d(i + 1)
            e2e_step(i)
        torch.cuda.synchronize()

    e2e_steps = max(4, min(steps, 20))
    e2e_run(4)
    barrier()
    t0 = time.perf_counter()
    s0, s1 = ev(), ev()
    s0.record()
    e2e_run(e2e_steps)
    s1.record()
    torch.cuda.synchronize()
    e_ms = torch.tensor([s0.elapsed_time(s1)], device=dev)
    if world > 1:
        dist.all_reduce(e_ms, op=dist.ReduceOp.MAX)
    barrier()
    e2e_wall = time.perf_counter() - t0
    e2e = {
        "value": world * T * e2e_steps / (float(e_ms.item()) / 1e3), "unit": "tokens/s",
        "h2d_bytes_per_step": dxs[0].numel() * 2 + dis[0].numel() * 8 + dws[0].numel() * 4, "d2h_bytes_per_step": douts[0].numel() * 2,
        "ms_per_step": float(e_ms.item()) / e2e_steps, "wall_s": e2e_wall, "steps": e2e_steps,
        "how": "Buffer.dispatch + Buffer.combine from pinned host buffers; H2D of the step's x/ids/weights and D2H of its output "
               "inside the timed region, double-buffered on copy streams; per-rank copies, aggregate tokens/s",
    }

    # ---- low-latency decode mode: CUDA-graph replay
    ll = None
    if not args.no_ll:
        try:
            ll = bench_ll(buf, rank, world, dev, barrier)
        except Exception as ex:  # keep the headline number even if the informative part fails
            ll = {"error": repr(ex)[:300]}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu, _ = cpu_baseline(1024, 5)

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "tokens/s", "n_gpus": world, "steps": steps, "warmup": warm,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": DTYPE, "data": "synthetic", "config": _config(world), "tokens_per_s_per_gpu": value / world,
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "clocks": clocks,
            # kernels of libb200ep.so inside the timed region: per step layout + dispatch + combine (EP=1), plus the
            # receive-side unpack kernel at EP>1
            "gpu_launches": steps * (3 if world == 1 else 4), "ll": ll,
            "host_sync_free": nosync, "grouped": grouped, "parity": parity_res,
        }
        os.write(real_stdout, (json.dumps(line) + "\n").encode())
    buf.destroy()
    if world > 1:
        dist.destroy_process_group()
    return 0


def bench_ll(buf, rank, world, dev, barrier, replays: int = 200):
    """Low-latency mode (BASELINE.json configs[2] and the 32-256 tokens/rank sweep of configs[4]):
    dispatch+combine captured in ONE CUDA graph and replayed; per-replay time from CUDA events, max
    over ranks of the p50.  The headline entry is 128 tokens/rank; `sweep` adds 32/64/256 and the
    EPLB shape (288 physical experts, 36 per rank at EP=8) when it divides over the ranks."""
    out = bench_ll_one(buf, rank, world, dev, barrier, LL_T, E, replays)
    sweep = {}
    for tok in (32, 64, 256):
        sweep[f"{tok}"] = bench_ll_one(buf, rank, world, dev, barrier, tok, E, 100, max_tok=256)["p50_us"]
    if 288 % world == 0:
        sweep["128_E288"] = bench_ll_one(buf, rank, world, dev, barrier, LL_T, 288, 100)["p50_us"]
    out["sweep_p50_us"] = sweep
    return out


def bench_ll_one(buf, rank, world, dev, barrier, LL_T, E, replays, max_tok=None):
    import torch.distributed as dist

    max_tok = max_tok or LL_T
    g0 = torch.Generator().manual_seed(99 + rank)
    x = (torch.randn(LL_T, H, generator=g0) * (torch.rand(LL_T, 1, generator=g0) * 4)).to(torch.bfloat16)
    vals, ids = torch.topk(torch.randn(LL_T, E, generator=g0), K, dim=-1, sorted=False)
    w = torch.softmax(vals, dim=-1).float().contiguous()
    ids = ids.to(torch.int64).contiguous()
    x, ids, w = x.to(dev), ids.to(dev), w.to(dev)
    el, cap = E // world, world * max_tok
    y = torch.randn(el, cap, H, device=dev).to(torch.bfloat16)
    out = torch.empty(LL_T, H, dtype=torch.bfloat16, device=dev)
    from b200ep import LlHandle
    from b200ep import _lib as L

    h = LlHandle(0, LL_T, H, K, E, max_tok, True)

    def one():
        buf.ll_dispatch_launch(h, x, ids, False, L.LL_PHASE_ALL)
        buf.low_latency_combine(y, ids, w, h, out, phases=L.PHASE_ALL)

    s = torch.cuda.Stream(device=dev)
    with torch.cuda.stream(s):
        for _ in range(3):
            one()
        s.synchronize()
        barrier()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s, capture_error_mode="relaxed"):
            one()
        barrier()
        for _ in range(10):
            g.replay()
        s.synchronize()
        barrier()
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(replays + 1)]
        evs[0].record(s)
        for i in range(replays):
            g.replay()
            evs[i + 1].record(s)
        s.synchronize()
    us = sorted(evs[i].elapsed_time(evs[i + 1]) * 1e3 for i in range(replays))
    p50, p99 = us[len(us) // 2], us[int(len(us) * 0.99) - 1]
    t = torch.tensor([p50, p99], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    barrier()
    # algorithmic floor: per (token, expert) message H + 4H/128 out, 2H back, (W-1)/W remote
    del y
    return {"tokens_per_rank": LL_T, "experts": E, "max_tokens_per_rank": max_tok, "mode": "fp8 dispatch + bf16 combine, one CUDA graph (2 kernels: count+send+wait, push+wait+reduce), replayed",
            "p50_us": float(t[0].item()), "p99_us": float(t[1].item()), "replays": replays, "ep_size": world}


if __name__ == "__main__":
    sys.exit(main())
