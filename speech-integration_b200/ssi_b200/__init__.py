"""ssi_b200 — B200-native (sm_100a) fused tied-LM-head + chunked cross-entropy for speech-integration.

Public surface (mirrors the reference's loss-module call signature, ssi/trainer.py:299-304,385-395):

* :class:`FusedLinearCrossEntropyLoss` — drop-in for torchtune's ``CEWithChunkedOutputLoss``
* :class:`HiddenPassthroughOutput` / :class:`TiedEmbeddingB200` / :func:`install` — model-side hooks
* :mod:`ssi_b200.ops` — tensor-level wrappers over the C ABI in ``include/ssi_b200.h``
* :mod:`ssi_b200.dist` — data-parallel helpers (all-reduce of the tied gradient and of the loss scalars)

Everything computes in ``libssi_b200.so`` (hand-written CUDA, tcgen05/TMEM/TMA); there is no CPU fallback.
"""
from . import _lib, build, ops  # noqa: F401
from .loss import FusedLinearCrossEntropyLoss, fused_linear_cross_entropy  # noqa: F401
from .model_hooks import HiddenPassthroughOutput, TiedEmbeddingB200, install  # noqa: F401

__all__ = [
    "FusedLinearCrossEntropyLoss",
    "fused_linear_cross_entropy",
    "HiddenPassthroughOutput",
    "TiedEmbeddingB200",
    "install",
    "ops",
    "build",
]
