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
        if weight.dtype == torch.bfloat16 and weight.is_contiguous():
            return ops.embedding_gather(weight, tokens)  # the library's own row gather
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


class _RMSNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, scale, eps):
        shape = x.shape
        x2 = x.reshape(-1, shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        need = x.requires_grad or scale.requires_grad
        y, rstd = ops.rmsnorm_fwd(x2, scale, eps, want_rstd=need)
        if need:
            ctx.save_for_backward(x2, scale, rstd)
        ctx.shape = shape
        return y.reshape(shape)

    @staticmethod
    def backward(ctx, dy):
        x2, scale, rstd = ctx.saved_tensors
        dy2 = dy.reshape(x2.shape)
        if dy2.dtype != torch.bfloat16:
            dy2 = dy2.to(torch.bfloat16)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx, dscale = ops.rmsnorm_bwd(dy2, x2, scale, rstd)
        return dx.reshape(ctx.shape), dscale.to(scale.dtype), None


class RMSNormB200(nn.Module):
    """Drop-in for torchtune's ``RMSNorm`` (tt/torchtune/modules/rms_norm.py:12-43: same ``scale`` parameter,
    ``eps`` and ``normalized_shape`` attributes, fp32 math, output in the input dtype) for the final norm that
    feeds the path (``model.norm``, tt/torchtune/modules/transformer.py:653).  bf16 CUDA tensors only."""

    def __init__(self, dim: int, eps: float = 1e-6, device=None, dtype=None) -> None:
        super().__init__()
        self.normalized_shape = (dim,)
        self.eps = eps
        self.scale = nn.Parameter(torch.ones(dim, device=device, dtype=dtype))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return _RMSNormFn.apply(x, self.scale, self.eps)

    @classmethod
    def from_norm(cls, norm: nn.Module) -> "RMSNormB200":
        new = cls(norm.scale.numel(), eps=norm.eps, device="meta", dtype=norm.scale.dtype)
        new.scale = norm.scale  # share the Parameter: optimizer / checkpoint key `norm.scale` unchanged
        return new


class FusedNormPassthrough(nn.Module):
    """Stands where ``model.norm`` stood when the final RMSNorm is fused into the loss: returns its input unchanged
    but still OWNS the ``scale`` Parameter (state-dict key ``norm.scale``, optimizer membership unchanged)."""

    def __init__(self, norm: nn.Module) -> None:
        super().__init__()
        self.normalized_shape = tuple(norm.scale.shape)
        self.eps = norm.eps
        self.scale = norm.scale

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return x


def install(model: nn.Module, num_output_chunks: int = 8, ignore_index: int = -100,
            replace_embedding: bool = True, mode: str | None = None,
            replace_norm: bool = False, fuse_norm: bool = False) -> FusedLinearCrossEntropyLoss:
    """Mirror of ``Trainer._setup_loss`` (ssi/trainer.py:299-304) for the fused path.

    Replaces ``model.output`` by a hidden passthrough, optionally swaps ``model.tok_embeddings`` for
    :class:`TiedEmbeddingB200` (sharing the same Parameter), calls ``model.set_num_output_chunks`` when the model
    has it, and returns the loss module to use as ``loss_fn``. ``mode``: ``"single_pass"`` (default) or
    ``"recompute"``, see :func:`ssi_b200.loss.fused_linear_cross_entropy`.  ``replace_norm``: also swap the final
    ``model.norm`` (tt/torchtune/modules/transformer.py:653) for :class:`RMSNormB200`, sharing its ``scale``.
    ``fuse_norm``: instead, fuse that norm into the loss (row gather / dX scatter of single-pass mode):
    ``model.norm`` becomes a :class:`FusedNormPassthrough` and the model hands pre-norm hidden states to the loss.
    """
    norm = getattr(model, "norm", None)
    fused_norm = None
    if fuse_norm:
        if norm is None or not (hasattr(norm, "scale") and hasattr(norm, "eps")):
            raise AttributeError("ssi_b200.install(fuse_norm=True): model.norm has no `scale` / `eps`")
        fused_norm = norm if isinstance(norm, FusedNormPassthrough) else FusedNormPassthrough(norm)
        model.norm = fused_norm
    elif (replace_norm and norm is not None and hasattr(norm, "scale") and hasattr(norm, "eps")
            and not isinstance(norm, RMSNormB200)):
        model.norm = RMSNormB200.from_norm(norm)
    emb = model.tok_embeddings
    if replace_embedding and isinstance(emb, nn.Embedding) and not isinstance(emb, TiedEmbeddingB200):
        emb = TiedEmbeddingB200.from_embedding(emb)
        model.tok_embeddings = emb
    model.output = HiddenPassthroughOutput(emb)
    cls = make_torchtune_subclass() or FusedLinearCrossEntropyLoss
    loss_fn = cls(emb, num_output_chunks=num_output_chunks, ignore_index=ignore_index, mode=mode)
    if fused_norm is not None:
        loss_fn.set_fused_norm(fused_norm)
    if hasattr(model, "set_num_output_chunks"):
        model.set_num_output_chunks(loss_fn.num_output_chunks)
    return loss_fn
