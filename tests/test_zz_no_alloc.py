"""b200ep_layout_noalloc: the routing layout for callers that cannot allocate device memory.

Who needs it: a serving worker that has already given the rest of the GPU to its KV-cache pool
(a cudaMalloc at that point fails or cuts into the pool), and small tools that want per-rank /
per-expert counts and send offsets for a batch of top-k ids without creating a b200ep context
(creating one allocates the multi-gigabyte symmetric region).  The entry point therefore takes no
context and keeps its kernel scratch in a fixed array inside the library; the caller still owns the
input and output buffers (torch tensors here).

The test calls it through the C ABI (ctypes) and compares every output with the existing entry
point b200ep_layout on the same ids -- integers, so the tolerance is "identical", as in the other
layout tests -- on two calls in a row (the second call finds the first call's leftovers in the
shared array), then checks that a batch beyond the fixed maximum is refused.  The file name sorts
last so that a `pytest -x` run has executed every other test before this one.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _call_noalloc(lib, world, ids, E):
    T, K = ids.shape
    dev = ids.device
    tpr = torch.full((world,), -7, dtype=torch.int32, device=dev)
    tpe = torch.full((E,), -7, dtype=torch.int32, device=dev)
    inr = torch.zeros((T, world), dtype=torch.bool, device=dev)
    off = torch.full((T, world), -7, dtype=torch.int32, device=dev)
    rc = lib.b200ep_layout_noalloc(world, ids.data_ptr(), int(ids.dtype == torch.int64), T, K, E, tpr.data_ptr(),
                                   tpe.data_ptr(), inr.data_ptr(), off.data_ptr(),
                                   torch.cuda.current_stream().cuda_stream)
    return rc, tpr, tpe, inr, off


def test_layout_noalloc_matches_layout_on_two_consecutive_calls():
    assert torch.cuda.is_available()
    from b200ep import Buffer, _lib

    lib = _lib.load()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(0)
    W, E, K = 8, 256, 8
    buf = Buffer(None, sym_bytes=8 << 20, rank=3, world=W, _connect=False)  # the existing entry point needs a context
    try:
        g = torch.Generator().manual_seed(2024)
        for call, (T, dtype) in enumerate([(4096, torch.int64), (1000, torch.int32)]):
            ids = torch.stack([torch.randperm(E, generator=g)[:K] for _ in range(T)]).to(dtype)
            ids[g.initial_seed() % T] = -1  # a token routed nowhere
            ids = ids.to(dev).contiguous()
            rc, tpr, tpe, inr, off = _call_noalloc(lib, W, ids, E)
            assert rc == 0, lib.b200ep_last_error()
            ref = buf.get_dispatch_layout(ids, E)
            torch.cuda.synchronize()
            assert torch.equal(tpr, ref.tokens_per_rank), f"call {call}: tokens_per_rank"
            assert torch.equal(tpe, ref.tokens_per_expert), f"call {call}: tokens_per_expert"
            assert torch.equal(inr, ref.is_token_in_rank), f"call {call}: is_token_in_rank"
            assert torch.equal(off, ref.send_off), f"call {call}: send_off"
    finally:
        buf.destroy()


def test_layout_noalloc_matches_the_cpu_oracle_on_a_side_stream():
    """Same entry point against oracle/layout.py directly (not only against the other CUDA entry point), issued on
    a stream of the caller's own, for one rank and for eight, with an int64 batch whose last 256-token block is
    partly empty."""
    import numpy as np

    from b200ep import _lib
    from oracle import layout as olayout

    lib = _lib.load()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(0)
    side = torch.cuda.Stream(device=dev)
    g = torch.Generator().manual_seed(7)
    for W, E, K, T in ((1, 64, 4, 777), (8, 288, 8, 2049)):
        ids = torch.stack([torch.randperm(E, generator=g)[:K] for _ in range(T)]).to(torch.int64)
        ids[torch.rand(T, K, generator=g) < 0.1] = -1
        want = olayout.dispatch_layout(ids.numpy(), E, W)
        ids = ids.to(dev).contiguous()
        torch.cuda.synchronize()
        with torch.cuda.stream(side):
            rc, tpr, tpe, inr, off = _call_noalloc(lib, W, ids, E)
            rc2, tpr2, tpe2, inr2, off2 = _call_noalloc(lib, W, ids, E)  # back to back on the same stream: serialised
        side.synchronize()
        assert rc == 0 and rc2 == 0, lib.b200ep_last_error()
        for got in ((tpr, tpe, inr, off), (tpr2, tpe2, inr2, off2)):
            assert np.array_equal(got[0].cpu().numpy(), want[0])
            assert np.array_equal(got[1].cpu().numpy(), want[1])
            assert np.array_equal(got[2].cpu().numpy().astype(np.uint8), want[2])
            assert np.array_equal(got[3].cpu().numpy(), want[3])


def test_layout_noalloc_rejects_batches_beyond_its_fixed_scratch():
    from b200ep import _lib

    lib = _lib.load()
    dev = torch.device("cuda:0")
    ids = torch.zeros((65537, 2), dtype=torch.int32, device=dev)
    rc, *_ = _call_noalloc(lib, 2, ids, 16)
    assert rc == 4  # B200EP_ERR_CAPACITY
    msg = lib.b200ep_last_error().decode()
    assert "65536" in msg and "b200ep_layout" in msg
    with pytest.raises(_lib.B200EPError):
        _lib.check(rc, "layout_noalloc")
    # the largest accepted batch still works
    rc, tpr, tpe, _, _ = _call_noalloc(lib, 2, ids[:65536].contiguous(), 16)
    torch.cuda.synchronize()
    assert rc == 0 and int(tpe[0]) == 2 * 65536 and tpr.tolist() == [65536, 0]
