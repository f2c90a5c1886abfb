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
        ids = tokens.reshape(-1)
        if not ids.is_contiguous():
            ids = ids.contiguous()
        # the radix sort of the ids that the backward's segmented reduction needs depends on the ids only: start it
        # here, so that nothing but the reduction itself is left for the tail of the step
        ctx.sort_ws = ops.embedding_grad_sort(ids, weight.size(0), weight.size(1))
        return ops.embedding_gather(weight, tokens)  # the library's own row gather

    @staticmethod
    def backward(ctx, grad_out):
        module = ctx.module
        w = module.weight
        D = grad_out.size(-1)
        dH = grad_out.reshape(-1, D)
        if dH.dtype != torch.bfloat16:
            dH = dH.to(torch.bfloat16)
        if not dH.is_contiguous():
            dH = dH.contiguous()
        sort_ws, ctx.sort_ws = ctx.sort_ws, None
        sync = getattr(module, "sync", None)
        if sync is not None:
            # data parallel: the LAST contribution of the micro-batch to the tied gradient (ssi_b200.sync)
            sync.embedding_backward(dH, sort_ws)
            return None, None, None
        if w.grad is None:
            w.grad = torch.zeros_like(w)
        fold = getattr(module, "fold", None)
        ps, fws = fold.sparse_args(dH.size(0)) if fold is not None else (None, None)
        ops.embedding_grad_scatter_sorted(dH, sort_ws, w.grad, post_scale=ps, fold_ws=fws)
        return None, None, None


class TiedEmbeddingB200(nn.Embedding):
    """``nn.Embedding`` (same parameter name ``weight``, same forward) with the B200 row gather forward and the
    sorted-id segmented scatter-add backward.  bf16 CUDA weights only (anything else is ``nn.Embedding``'s job)."""

    sync = None  # ssi_b200.sync.TiedGradSync when install() was given a process group
    fold = None  # ssi_b200.fold.TiedGradFold when install(fold_scale_grads=True)

    def forward(self, tokens: torch.Tensor) -> torch.Tensor:
        w = self.weight
        if w.is_cuda and w.dtype == torch.bfloat16 and w.is_contiguous() and self.max_norm is None:
            if w.requires_grad and torch.is_grad_enabled():
                return _EmbeddingFn.apply(w, tokens, self)
            return ops.embedding_gather(w.detach(), tokens)  # eval / frozen: the same gather, nothing saved
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


def attach_fold(loss_fn: FusedLinearCrossEntropyLoss, emb: nn.Module, enabled: bool = True):
    """Give the loss and the tied embedding a shared :class:`ssi_b200.fold.TiedGradFold` (or take it away)."""
    sync = getattr(loss_fn, "sync", None)
    fold = None
    if enabled:
        if loss_fn.mode != "single_pass" or not isinstance(emb, TiedEmbeddingB200):
            raise ValueError("ssi_b200.install(fold_scale_grads=True): needs mode='single_pass' and "
                             "replace_embedding=True (the fold lives in apply_dw / the peer reduce and in the "
                             "embedding backward)")
        if sync is not None and sync.peer is None:
            raise ValueError("ssi_b200.install(fold_scale_grads=True): exchange='nccl' all-reduces bf16 gradients "
                             "after the step; there is no kernel of the library left to fold into")
        from .fold import TiedGradFold

        fold = TiedGradFold(emb.weight)
    object.__setattr__(loss_fn, "fold", fold)
    object.__setattr__(emb, "fold", fold)
    if sync is not None:
        sync.fold = fold
    return fold


def install(model: nn.Module, num_output_chunks: int = 8, ignore_index: int = -100,
            replace_embedding: bool = True, mode: str | None = None,
            replace_norm: bool = False, fuse_norm: bool = False, process_group=None,
            exchange: str | None = None, max_tokens_per_rank: int = 65536,
            fold_scale_grads: bool = False) -> FusedLinearCrossEntropyLoss:
    """Mirror of ``Trainer._setup_loss`` (ssi/trainer.py:299-304) for the fused path.

    Replaces ``model.output`` by a hidden passthrough, optionally swaps ``model.tok_embeddings`` for
    :class:`TiedEmbeddingB200` (sharing the same Parameter), calls ``model.set_num_output_chunks`` when the model
    has it, and returns the loss module to use as ``loss_fn``. ``mode``: ``"single_pass"`` (default) or
    ``"recompute"``, see :func:`ssi_b200.loss.fused_linear_cross_entropy`.  ``replace_norm``: also swap the final
    ``model.norm`` (tt/torchtune/modules/transformer.py:653) for :class:`RMSNormB200`, sharing its ``scale``.
    ``fuse_norm``: instead, fuse that norm into the loss (row gather / dX scatter of single-pass mode):
    ``model.norm`` becomes a :class:`FusedNormPassthrough` and the model hands pre-norm hidden states to the loss.

    ``process_group`` (one process per GPU; the reference's data-level scaffold is ``DistributedSampler``,
    ssi/data/__init__.py:82-83): the tied gradient — and nothing else — is summed over the group's ranks, in the order
    a training step produces its parts (:mod:`ssi_b200.sync`; semantics of tt/recipes/full_finetune_distributed.py:
    844-851: losses are this rank's share of the global mean, gradients those of the global mean).  ``exchange``:
    ``"peer"`` (default: the library's own NVLink peer-memory kernels), ``"nvls"`` (the same with the dense reduce done
    inside the NVSwitch, ``multimem.ld_reduce``; buffers from torch's symmetric-memory allocator) or ``"nccl"`` (bf16
    all-reduce baseline);
    ``max_tokens_per_rank`` sizes the compact-row area of the peer exchange (most token rows of one micro-batch).

    ``fold_scale_grads``: give the loss a :class:`ssi_b200.fold.TiedGradFold` (``loss_fn.fold``): the trainer's
    ``scale_grads`` and the tied weight's share of ``clip_grad_norm_`` (ssi/trainer.py:404-408) ride along in the kernels
    that write the tied gradient anyway (single-pass mode; with a process group: the peer exchange)."""
    norm = getattr(model, "norm", None)
    # what uninstall() puts back (kept outside nn.Module's registries: no state-dict keys, no extra parameters);
    # a second install() on the same model keeps the ORIGINAL parts of the first
    saved = getattr(model, "_ssi_b200_saved", None) or {
        "output": getattr(model, "output", None), "tok_embeddings": model.tok_embeddings, "norm": norm,
        "has_norm": hasattr(model, "norm"), "num_output_chunks": getattr(model, "num_output_chunks", None)}
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
    if process_group is not None:
        if not isinstance(emb, TiedEmbeddingB200):
            raise ValueError("ssi_b200.install(process_group=...): needs replace_embedding=True (the embedding "
                             "backward is where a micro-batch's exchange of the tied gradient is completed)")
        from .sync import TiedGradSync

        if loss_fn.mode != "single_pass":
            # the peer exchange sums the fp32 dW_acc that only single-pass mode forms
            if exchange in ("peer", "nvls"):
                raise ValueError(f"ssi_b200: exchange={exchange!r} needs mode='single_pass'")
            exchange = exchange or "nccl"
        sync = TiedGradSync(emb.weight, process_group, exchange=exchange, max_rows=max_tokens_per_rank)
        object.__setattr__(loss_fn, "sync", sync)
        object.__setattr__(emb, "sync", sync)
    if fold_scale_grads:
        attach_fold(loss_fn, emb)
    saved["loss_fn"] = loss_fn
    object.__setattr__(model, "_ssi_b200_saved", saved)
    return loss_fn


def uninstall(model: nn.Module) -> None:
    """Undo :func:`install`: ``model.output`` (the ``TiedLinear`` of tt/torchtune/models/llama3_2/
    _component_builders.py:123), ``model.tok_embeddings``, ``model.norm`` and the chunk count are the objects the model
    had before, so ``TransformerDecoder.forward`` returns LOGITS again (tt/torchtune/modules/transformer.py:651-661)
    -- for generation, for a caller that wants ``model.unembed``, or to hand the model back to the reference's own
    loss.  Parameters were shared all along (same ``tok_embeddings.weight`` / ``norm.scale`` objects), so weights,
    gradients and optimizer state carry over unchanged.  With a process group this also releases the exchange buffers
    (collective: every rank must call it).  No-op on a model :func:`install` never touched."""
    saved = getattr(model, "_ssi_b200_saved", None)
    if saved is None:
        return
    sync = getattr(saved.get("loss_fn"), "sync", None)
    if sync is not None:
        sync.close()
    model.tok_embeddings = saved["tok_embeddings"]
    model.output = saved["output"]
    if saved["has_norm"]:
        model.norm = saved["norm"]
    if saved["num_output_chunks"] is not None and hasattr(model, "set_num_output_chunks"):
        model.set_num_output_chunks(saved["num_output_chunks"])
    object.__delattr__(model, "_ssi_b200_saved")
