"""ssi_b200 — B200-native (sm_100a) fused tied-LM-head + chunked cross-entropy for speech-integration.

Public surface (mirrors the reference's loss-module call signature, ssi/trainer.py:299-304,385-395):

* :class:`FusedLinearCrossEntropyLoss` — drop-in for torchtune's ``CEWithChunkedOutputLoss``
* :class:`HiddenPassthroughOutput` / :class:`TiedEmbeddingB200` / :func:`install` — model-side hooks
* :mod:`ssi_b200.ops` — tensor-level wrappers over the C ABI in ``include/ssi_b200.h``
* :class:`RMSNormB200` — the final norm in front of the path
* :mod:`ssi_b200.sync` / :mod:`ssi_b200.peer` / :mod:`ssi_b200.dist` — data parallelism behind ``install(model,
  process_group=...)``: token counts, LM-head part and embedding part of the tied gradient exchanged over NVLink peer
  memory (or NCCL) in the order a training step produces them

Everything computes in ``libssi_b200.so`` (hand-written CUDA, tcgen05/TMEM/TMA); there is no CPU fallback.
"""
from . import _lib, build, ops  # noqa: F401
from .loss import FusedLinearCrossEntropyLoss, fused_linear_cross_entropy  # noqa: F401
from .model_hooks import HiddenPassthroughOutput, RMSNormB200, TiedEmbeddingB200, install, uninstall  # noqa: F401

__all__ = [
    "FusedLinearCrossEntropyLoss",
    "fused_linear_cross_entropy",
    "HiddenPassthroughOutput",
    "TiedEmbeddingB200",
    "RMSNormB200",
    "install",
    "uninstall",
    "ops",
    "build",
]
