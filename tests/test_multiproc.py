"""Multi-process paths: world_size-2 gloo on the CPU (host logic), torchrun over CUDA IPC on GPUs."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, "tests", "mp_worker.py")


def _torchrun(n, args, port, timeout=300):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", str(port)] + args
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT)


@pytest.mark.parametrize("world", [2, 4])
def test_gloo_world_host_logic(world):
    r = _torchrun(world, [WORKER, "cpu"], 29611 + world)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    for k in range(world):
        assert f"CPU-MP-OK rank {k}/{world}" in r.stdout


def test_bench_reference_arm_under_torchrun():
    """--impl reference: rank 0 alone runs and prints one JSON line, the other ranks exit 0."""
    r = _torchrun(2, [os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                      "--warmup", "0"], 29631, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    j = json.loads(lines[0])
    assert j["impl"] == "reference" and j["value"] > 0 and j["cpu_baseline"]["kind"] == "port"
    assert j["e2e"]["h2d_bytes_per_step"] == 0


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4, 8])
def test_gpu_ipc_world(world):
    import torch

    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    r = _torchrun(world, [WORKER, "gpu"], 29651 + world, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-6000:]
    for k in range(world):
        assert f"GPU-MP-OK rank {k}/{world}" in r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4, 8])
def test_gpu_ipc_world_baseline_shapes(world):
    import torch

    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    r = _torchrun(world, [WORKER, "gpu_full"], 29671 + world, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-6000:]
    for k in range(world):
        assert f"GPU-MP-FULL-OK rank {k}/{world}" in r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("backend", ["deepep_high_throughput", "deepep_low_latency"])
@pytest.mark.parametrize("world", [2, 8])
def test_gpu_vllm_managers_world(world, backend):
    """vLLM builds its own DeepEP all2all manager for the llm-d backend names at data_parallel_size = world, the manager's
    handle is the b200ep shim, and vLLM's own prepare/finalize classes give oracle-exact results on it with two
    micro-batches in flight (tests/mp_vllm_worker.py)."""
    import torch

    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    pytest.importorskip("vllm")
    r = _torchrun(world, [os.path.join(ROOT, "tests", "mp_vllm_worker.py"), backend], 29691 + world, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-8000:]
    for k in range(world):
        assert f"VLLM-MGR-OK {backend} rank {k}/{world}" in r.stdout
