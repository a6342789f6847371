"""pytest configuration: the ``gpu`` marker and import paths.

``-m "not gpu"`` tests run in the GPU-less build container (oracle vs golden
vectors, host logic, C-ABI symbols); ``-m gpu`` tests are the parity tests
proper and call the CUDA kernels through the C-ABI on a B200.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "llm-d_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
