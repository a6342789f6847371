"""GPU parity: layout kernel and FP8 quantiser vs the CPU oracle, through the C ABI."""
import numpy as np
import pytest
import torch

from _util import bf16_bits, make_rank_inputs, u8
from oracle import fp8, layout

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch.device("cuda:0")


@pytest.mark.parametrize("W", [1, 2, 4, 8])
@pytest.mark.parametrize("T,K,E,idt,mask", [
    (0, 8, 256, torch.int64, 0.0), (1, 8, 256, torch.int64, 0.0), (127, 8, 256, torch.int32, 0.0),
    (128, 8, 32, torch.int64, 0.3), (4096, 8, 256, torch.int64, 0.0), (4096, 8, 288, torch.int32, 0.0),
    (5000, 6, 128, torch.int64, 0.1)])
def test_layout_bit_exact(dev, W, T, K, E, idt, mask):
    if E % W:
        pytest.skip("experts must divide over ranks")
    from b200ep import Buffer

    buf = Buffer(None, sym_bytes=8 << 20, rank=W - 1, world=W, _connect=False)
    try:
        for rep, dist in enumerate(["uniform", "one_rank", "zipf"]):
            _, ids, _ = make_rank_inputs(T, 128, E, K, seed=T + rep, dist=dist, world=W, id_dtype=idt, mask_frac=mask)
            lay = buf.get_dispatch_layout(ids.to(dev), E)
            tpr, tpe, inr, off = layout.dispatch_layout(ids.numpy(), E, W)
            assert np.array_equal(lay.tokens_per_rank.cpu().numpy(), tpr)
            assert np.array_equal(lay.tokens_per_expert.cpu().numpy(), tpe)
            if T:
                assert np.array_equal(lay.is_token_in_rank.cpu().numpy().astype(np.uint8), inr)
                assert np.array_equal(lay.send_off.cpu().numpy(), off)
    finally:
        buf.destroy()


def test_layout_many_tokens_multi_tile(dev):
    """More than 1024 blocks' worth of tokens: the multi-tile path (and a size-independent check)."""
    from b200ep import Buffer

    W, E, K, T = 8, 256, 8, 1024 * 256 + 777
    buf = Buffer(None, sym_bytes=8 << 20, rank=0, world=W, _connect=False)
    try:
        g = torch.Generator().manual_seed(1)
        ids = torch.randint(0, E, (T, K), generator=g, dtype=torch.int64)
        lay = buf.get_dispatch_layout(ids.to(dev), E)
        tpr, tpe, inr, off = layout.dispatch_layout(ids.numpy(), E, W)
        assert np.array_equal(lay.tokens_per_rank.cpu().numpy(), tpr)
        assert np.array_equal(lay.tokens_per_expert.cpu().numpy(), tpe)
        assert np.array_equal(lay.send_off.cpu().numpy(), off)
    finally:
        buf.destroy()


def test_layout_replays_under_a_cuda_graph(dev):
    """The kernel keeps its call state on the device (tickets, epoch, histogram) and rewinds it itself: a captured
    launch must give the right answer on every replay, with different ids each time."""
    from b200ep import Buffer, _lib

    W, E, K, T = 4, 64, 8, 3000
    buf = Buffer(None, sym_bytes=8 << 20, rank=1, world=W, _connect=False)
    try:
        ids = torch.zeros(T, K, dtype=torch.int64, device=dev)
        tpr = torch.empty(W, dtype=torch.int32, device=dev)
        tpe = torch.empty(E, dtype=torch.int32, device=dev)
        inr = torch.empty((T, W), dtype=torch.bool, device=dev)
        off = torch.empty((T, W), dtype=torch.int32, device=dev)
        s = torch.cuda.Stream(device=dev)

        def call():
            _lib.check(buf.lib.b200ep_layout(buf.ctx, ids.data_ptr(), 1, T, K, E, tpr.data_ptr(), tpe.data_ptr(),
                                             inr.data_ptr(), off.data_ptr(), torch.cuda.current_stream().cuda_stream))

        with torch.cuda.stream(s):
            call()
            s.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=s):
                call()
            for rep in range(3):
                _, new_ids, _ = make_rank_inputs(T, 128, E, K, seed=50 + rep, dist=("uniform", "zipf", "one_rank")[rep], world=W,
                                                 mask_frac=0.1 * rep)
                ids.copy_(new_ids.to(dev))
                g.replay()
                s.synchronize()
                want = layout.dispatch_layout(new_ids.numpy(), E, W)
                assert np.array_equal(tpr.cpu().numpy(), want[0]) and np.array_equal(tpe.cpu().numpy(), want[1])
                assert np.array_equal(inr.cpu().numpy().astype(np.uint8), want[2])
                assert np.array_equal(off.cpu().numpy(), want[3])
    finally:
        buf.destroy()


def _adversarial(H):
    rows = []
    rows.append(np.zeros(H, np.float32))                       # all-zero groups -> eps path
    r = np.zeros(H, np.float32); r[::128] = 448.0; rows.append(r)   # amax == 448 -> scale exactly 1
    r = np.full(H, 2.0**-20, np.float32); rows.append(r)       # tiny
    r = np.full(H, 3.0e38, np.float32); rows.append(r)         # huge (bf16 max range)
    r = np.linspace(-1, 1, H).astype(np.float32); r[7] = 1.75; rows.append(r)  # quotient at a power of two
    # values at e4m3 rounding midpoints relative to amax = 448 (scale 1)
    mids = [(1 + (m + 0.5) / 8) * 2.0**e for e in range(-6, 8) for m in range(8)]
    r = np.resize(np.array(mids, np.float32), H); r[0] = 448.0; r[128::128] = 448.0; rows.append(r)
    return np.stack(rows)


@pytest.mark.parametrize("H", [128, 1024, 2048, 7168])
@pytest.mark.parametrize("ue8m0", [False, True])
def test_quant_bit_exact_vs_oracle(dev, H, ue8m0):
    from b200ep import Buffer

    x, _, _ = make_rank_inputs(333, H, 8, 2, seed=H)
    adv = torch.from_numpy(_adversarial(H)).to(torch.bfloat16)
    x = torch.cat([x, adv, -adv], 0).contiguous()
    q, s = Buffer.quant_fp8(x.to(dev), ue8m0=ue8m0)
    oq, os_ = fp8.group_quant_fp8(bf16_bits(x), formula="div", ue8m0=ue8m0)
    assert np.array_equal(s.cpu().numpy().view(np.uint32), os_.view(np.uint32))
    assert np.array_equal(u8(q), oq)


def test_quant_matches_vllm_cuda_op(dev):
    """Pins the FP8 arithmetic against the reference's real quantiser: vLLM's CUDA op
    per_token_group_quant_8bit, reached through per_token_group_quant_fp8 (fp8_utils.py:510-591)."""
    try:
        from vllm.model_executor.layers.quantization.utils.fp8_utils import per_token_group_quant_fp8
    except Exception as e:  # pragma: no cover
        pytest.skip(f"vLLM quantiser not importable here: {e}")
    from b200ep import Buffer

    x, _, _ = make_rank_inputs(1024, 7168, 8, 2, seed=5)
    adv = torch.from_numpy(_adversarial(7168)).to(torch.bfloat16)
    x = torch.cat([x, adv, -adv], 0).contiguous().to(dev)
    for ue in (False, True):
        rq, rs = per_token_group_quant_fp8(x, 128, use_ue8m0=ue)
        q, s = Buffer.quant_fp8(x, ue8m0=ue)
        bad = (rs.contiguous().view(torch.int32) != s.view(torch.int32)).nonzero()
        assert bad.numel() == 0, (f"scales differ (ue8m0={ue}) at {bad[:5].tolist()}: "
                                  f"vllm {[rs[i, j].item() for i, j in bad[:5].tolist()]} "
                                  f"ours {[s[i, j].item() for i, j in bad[:5].tolist()]}")
        assert torch.equal(rq.view(torch.uint8), q.view(torch.uint8)), f"fp8 bytes differ (ue8m0={ue})"
