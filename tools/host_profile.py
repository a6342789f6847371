"""Where does the host time of one EP=1 step go?  (cProfile over the public calls bench.py times.)
Not a test and not part of the product: a measurement aid, run on the GPU box as
  python tools/host_profile.py [steps]
"""
import cProfile
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "llm-d_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402

import b200ep  # noqa: E402
from bench import E, H, K, T, make_inputs  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 300
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
buf = b200ep.Buffer(None, sym_bytes=1 << 30, device=0, rank=0, world=1)
x, ids, w = (t.to(dev) for t in make_inputs(1))
y = torch.randn(T, H, device=dev).to(torch.bfloat16)
out = torch.empty(T, H, dtype=torch.bfloat16, device=dev)


def step(nosync):
    rx, rs, ridx, rw, pe, h = buf.dispatch(x, ids, w, E, quantize=True, num_worst_tokens=(T if nosync else 0))
    buf.combine(y[: h.num_recv], h, out)


for mode in (False, True):
    for _ in range(20):
        step(mode)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        step(mode)
    t_host = time.perf_counter() - t0
    torch.cuda.synchronize()
    t_all = time.perf_counter() - t0
    print(f"nosync={mode}: host loop {t_host / steps * 1e6:.1f} us/step, with final sync {t_all / steps * 1e6:.1f} us/step")
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(steps):
        step(mode)
    pr.disable()
    torch.cuda.synchronize()
    pstats.Stats(pr).sort_stats("tottime").print_stats(14)
buf.destroy()
