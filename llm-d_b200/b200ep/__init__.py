"""b200ep -- B200-native wide-EP MoE token dispatch/combine (host package).

Python/PyTorch above the C ABI of ``libb200ep.so`` (``include/b200ep.h``):
``Buffer`` is the per-rank endpoint, ``LocalWorld`` emulates several ranks on one
GPU for tests.  The vLLM-facing classes live in ``b200ep.vllm_backend`` and the
DeepEP-compatible surface in the sibling package ``deep_ep``.
"""
from ._lib import (B200EPError, LL_PHASE_ALL, LL_PHASE_COUNT, LL_PHASE_RECV, LL_PHASE_SEND, PHASE_ALL, PHASE_RECV,
                   PHASE_SEND, X_BF16, X_BF16_QUANT, X_FP8, load)
from .buffer import Buffer, HtHandle, Layout, LlHandle, LocalWorld

__all__ = ["Buffer", "LocalWorld", "Layout", "HtHandle", "LlHandle", "B200EPError", "load", "X_BF16", "X_BF16_QUANT",
           "X_FP8", "PHASE_SEND", "PHASE_RECV", "PHASE_ALL", "LL_PHASE_COUNT", "LL_PHASE_SEND", "LL_PHASE_RECV",
           "LL_PHASE_ALL"]
