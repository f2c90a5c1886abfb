"""Model-side hooks that route the reference model's output projection and embedding gradient through ssi_b200.

* :class:`HiddenPassthroughOutput` replaces ``TransformerDecoder.output`` (an arbitrary callable stored at
  tt/torchtune/modules/transformer.py:384 and invoked at :500 / :659, normally a ``TiedLinear``,
  tt/torchtune/modules/tied_linear.py:28-64): it returns the normed hidden chunk unchanged so that the model
  hands hidden states to the loss, which then fuses projection + CE.
* :class:`TiedEmbeddingB200` is an ``nn.Embedding`` whose backward is the sorted-id segmented scatter-add
  (``ssi_b200_embedding_grad_scatter``) into the SAME ``.grad`` buffer the LM-head gradient is written to.
* :func:`install` swaps both into a model that has ``tok_embeddings`` / ``output`` attributes
  (tt/torchtune/models/llama3_2/_component_builders.py:114-123) and returns the fused loss.
"""
from __future__ import annotations

import torch
from torch import nn

from . import ops
from .loss import FusedLinearCrossEntropyLoss, make_torchtune_subclass


class HiddenPassthroughOutput:
    """Callable with TiedLinear's shape of interface (``__call__(x)``, ``.tied_module``); not an nn.Module, so it
    adds no state-dict keys, exactly like TiedLinear (tt/torchtune/modules/tied_linear.py:28)."""

    def __init__(self, tied_module: nn.Module):
        if not hasattr(tied_module, "weight"):
            raise AttributeError(
                "Provided module does not have attribute 'weight'. Please check your tied_module."
            )
        self.tied_module = tied_module

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        return x


class _EmbeddingFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, weight, tokens, module):
        ctx.module = module
        ctx.save_for_backward(tokens)
        ctx.wshape = weight.shape
        return torch.nn.functional.embedding(tokens, weight)

    @staticmethod
    def backward(ctx, grad_out):
        (tokens,) = ctx.saved_tensors
        module = ctx.module
        w = module.weight
        D = grad_out.size(-1)
        dH = grad_out.reshape(-1, D)
        if dH.dtype != torch.bfloat16:
            dH = dH.to(torch.bfloat16)
        if not dH.is_contiguous():
            dH = dH.contiguous()
        ids = tokens.reshape(-1)
        if not ids.is_contiguous():
            ids = ids.contiguous()
        if w.grad is None:
            w.grad = torch.zeros_like(w)
        ops.embedding_grad_scatter(dH, ids, w.grad)
        return None, None, None


class TiedEmbeddingB200(nn.Embedding):
    """``nn.Embedding`` (same parameter name ``weight``, same forward) with the B200 scatter-add backward."""

    def forward(self, tokens: torch.Tensor) -> torch.Tensor:
        if self.weight.requires_grad and torch.is_grad_enabled() and self.weight.is_cuda:
            return _EmbeddingFn.apply(self.weight, tokens, self)
        return super().forward(tokens)

    @classmethod
    def from_embedding(cls, emb: nn.Embedding) -> "TiedEmbeddingB200":
        new = cls(emb.num_embeddings, emb.embedding_dim, padding_idx=emb.padding_idx, max_norm=emb.max_norm,
                  norm_type=emb.norm_type, scale_grad_by_freq=emb.scale_grad_by_freq, sparse=emb.sparse,
                  device="meta", dtype=emb.weight.dtype)
        new.weight = emb.weight  # share the Parameter: optimizer / checkpoint keys unchanged
        return new


def install(model: nn.Module, num_output_chunks: int = 8, ignore_index: int = -100,
            replace_embedding: bool = True, mode: str | None = None) -> FusedLinearCrossEntropyLoss:
    """Mirror of ``Trainer._setup_loss`` (ssi/trainer.py:299-304) for the fused path.

    Replaces ``model.output`` by a hidden passthrough, optionally swaps ``model.tok_embeddings`` for
    :class:`TiedEmbeddingB200` (sharing the same Parameter), calls ``model.set_num_output_chunks`` when the model
    has it, and returns the loss module to use as ``loss_fn``. ``mode``: ``"single_pass"`` (default) or
    ``"recompute"``, see :func:`ssi_b200.loss.fused_linear_cross_entropy`.
    """
    emb = model.tok_embeddings
    if replace_embedding and isinstance(emb, nn.Embedding) and not isinstance(emb, TiedEmbeddingB200):
        emb = TiedEmbeddingB200.from_embedding(emb)
        model.tok_embeddings = emb
    model.output = HiddenPassthroughOutput(emb)
    cls = make_torchtune_subclass() or FusedLinearCrossEntropyLoss
    loss_fn = cls(emb, num_output_chunks=num_output_chunks, ignore_index=ignore_index, mode=mode)
    if hasattr(model, "set_num_output_chunks"):
        model.set_num_output_chunks(loss_fn.num_output_chunks)
    return loss_fn
