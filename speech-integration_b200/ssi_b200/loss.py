"""Drop-in replacement for torchtune's ``CEWithChunkedOutputLoss`` that fuses the tied output projection.

Reference call signature (ssi/trainer.py:299-304, ssi/loss.py:7-22): the trainer builds the loss, reads
``num_output_chunks`` / ``ignore_index`` and calls ``loss_fn(model_output, labels)`` where ``model_output`` is a
list of ``num_output_chunks`` tensors ``[b, s/n, *]`` (labels ``[b, s]``) or a single 2-D tensor (labels
``[b*s]``).  With :class:`ssi_b200.HiddenPassthroughOutput` installed as ``model.output`` the model returns
the final normed HIDDEN states in those positions and this module does projection + CE in one fused pass.
"""
from __future__ import annotations

import os
from typing import Sequence, Union

import torch
from torch import nn

from . import ops


class _FusedLinearCE(torch.autograd.Function):
    """Recompute mode.  loss = mean (or sum) over valid rows of CE(X W^T, labels); saves only X, W, labels and the
    per-row logsumexp; backward recomputes the logits tiles."""

    @staticmethod
    def forward(ctx, X, W, labels, ignore_index, holder, n_valid_override, reduction="mean"):
        loss_out, lse, n_valid = ops.lmhead_ce_fwd(X, W, labels, ignore_index)
        ctx.ignore_index = ignore_index
        ctx.holder = holder
        if reduction == "sum":
            n_div = _ones_i64(X.device)
        else:
            n_div = n_valid if n_valid_override is None else n_valid_override
        ctx.save_for_backward(X, W, labels, lse, n_div)
        if holder is not None:
            holder.last_loss_sum = loss_out[0:1]
            holder.last_n_valid = n_valid
        if reduction == "sum":
            return loss_out[0]
        if n_valid_override is None:
            return loss_out[1]
        # data parallel: mean over the GLOBAL number of valid tokens (this rank's share of it)
        return loss_out[0] / n_valid_override[0].to(torch.float32)

    @staticmethod
    def backward(ctx, grad_out):
        X, W, labels, lse, n_div = ctx.saved_tensors
        g = grad_out.detach().to(torch.float32).reshape(1).contiguous()
        holder = ctx.holder
        wparam = holder.weight_param() if holder is not None else None
        write_into_grad = (
            wparam is not None and holder.accumulate_into_grad and wparam.data_ptr() == W.data_ptr()
        )
        sync = getattr(holder, "sync", None) if write_into_grad else None
        if sync is not None:
            # data parallel: this micro-batch's dW goes where it can be all-reduced alone (ssi_b200.sync)
            target, acc = sync.recompute_target()
            dX, _ = ops.lmhead_ce_bwd(X, W, labels, lse, g, n_div, ctx.ignore_index, dW=target, accumulate_dW=acc)
            return dX, None, None, None, None, None, None
        if write_into_grad:
            # Write the LM-head gradient straight into the tied weight's .grad buffer (the buffer the
            # embedding scatter-add and the DP all-reduce also use) and hand autograd nothing to accumulate.
            if wparam.grad is None:
                wparam.grad = torch.empty_like(wparam)
                dX, _ = ops.lmhead_ce_bwd(X, W, labels, lse, g, n_div, ctx.ignore_index, dW=wparam.grad,
                                          accumulate_dW=False)
            else:
                dX, _ = ops.lmhead_ce_bwd(X, W, labels, lse, g, n_div, ctx.ignore_index, dW=wparam.grad,
                                          accumulate_dW=True)
            return dX, None, None, None, None, None, None
        dX, dW = ops.lmhead_ce_bwd(X, W, labels, lse, g, n_div, ctx.ignore_index)
        return dX, dW, None, None, None, None, None


_CONST: dict = {}


def _one(device: torch.device) -> torch.Tensor:
    t = _CONST.get(("f", device))
    if t is None:
        t = torch.ones(1, dtype=torch.float32, device=device)
        _CONST[("f", device)] = t
    return t


def _ones_i64(device: torch.device) -> torch.Tensor:
    t = _CONST.get(("i", device))
    if t is None:
        t = torch.ones(1, dtype=torch.int64, device=device)
        _CONST[("i", device)] = t
    return t


class _FusedLinearCESinglePass(torch.autograd.Function):
    """Single-pass mode: when a gradient is needed, forward also forms dX and the fp32 dW sum (each logits tile
    is computed once; a bounded bf16 block of logits/dlogits is the only scratch). Backward only rescales by
    ``grad_out / grad_hint`` (1 when the trainer announced its multiplier) and rounds dW once into ``.grad``."""

    @staticmethod
    def forward(ctx, X, W, labels, ignore_index, holder, n_valid_override, grad_hint, grad_enabled=True,
                norm_scale=None, norm_eps=1e-6, reduction="mean"):
        ctx.holder = holder
        ctx.dscale = None
        ctx.sync = None
        # `grad_enabled` is sampled by the caller: inside Function.forward grad mode is always off, and under
        # torch.no_grad() / inference_mode() (the eval path, ssi/eval.py:27-33) needs_input_grad still reports the
        # Parameter -- forming gradients there would triple the work of a forward-only call
        ctx.with_grad = bool(grad_enabled and (ctx.needs_input_grad[0] or ctx.needs_input_grad[1]
                                               or (norm_scale is not None and ctx.needs_input_grad[8])))
        gcounts = None
        if ctx.with_grad:
            hint = _one(X.device) if grad_hint is None else grad_hint.detach().to(torch.float32).reshape(1)
            announced = grad_hint is not None
            kw = {}
            sync = getattr(holder, "sync", None)
            if sync is not None and reduction == "mean" and W.data_ptr() == holder.tied_module.weight.data_ptr():
                # Data parallel (ssi_b200.sync): the local token counts are summed over ranks beside the first logits
                # contraction; the divisor of the mean -- and the announced multiplier when it is "the global token
                # count" -- are device scalars that the first softmax-grad pass reads.
                local = getattr(holder, "local_counts", None)
                if local is None:
                    local = ops.count_valid(labels, ignore_index)
                gcounts, ghint, ev = sync.global_counts(local)
                n_valid_override = gcounts[1:2]
                if getattr(holder, "hint_is_global_count", False):
                    hint, announced = ghint, True
                kw["pre_grad_event"] = ev
                if sync.peer is not None:
                    sync.before_dw_acc_write()  # a dense exchange still reading the buffer (2nd loss of a micro-batch)
                    kw["dW_acc"], kw["dw_ready_event"] = sync.dw_acc, sync.dw_ready
                ctx.sync = sync
            nvt = _ones_i64(X.device) if reduction == "sum" else n_valid_override
            out = ops.lmhead_ce_fwd_grad(
                X, W, labels, hint, nvt, ignore_index,
                norm_scale=norm_scale.detach() if norm_scale is not None else None, norm_eps=norm_eps, **kw)
            loss_out, lse, n_valid, dX, dW_acc = out[:5]
            if norm_scale is not None:  # X was the PRE-norm state: dX is its gradient, dscale the scale's
                ctx.dscale, ctx.scale_dtype = out[5], norm_scale.dtype
            ctx.hint, ctx.dX, ctx.dW_acc, ctx.w_ptr = hint, dX, dW_acc, W.data_ptr()
            if ctx.sync is not None:
                ctx.sync.after_forward(hint, announced, X.size(0))
        else:
            if norm_scale is not None:  # forward only (eval): the stand-alone norm kernel, then K1
                X = ops.rmsnorm_fwd(X, norm_scale.detach(), norm_eps, want_rstd=False)[0]
            loss_out, lse, n_valid = ops.lmhead_ce_fwd(X, W, labels, ignore_index)
        if holder is not None:
            holder.last_loss_sum = loss_out[0:1]
            holder.last_n_valid = n_valid
            holder.last_global_counts = gcounts
        if reduction == "sum":
            return loss_out[0]
        if n_valid_override is None:
            return loss_out[1]
        return loss_out[0] / n_valid_override[0].to(torch.float32)

    @staticmethod
    def backward(ctx, grad_out):
        if ctx.dW_acc is None:
            raise RuntimeError("ssi_b200: single-pass backward called twice (the stashed gradients were consumed)")
        g = grad_out.detach().to(torch.float32).reshape(1).contiguous()
        holder = ctx.holder
        dX, dW_acc, hint = ctx.dX, ctx.dW_acc, ctx.hint
        ctx.dX = ctx.dW_acc = None
        dscale = None
        if ctx.dscale is not None:  # gradient of the fused norm's scale, formed for `hint` like dX and dW_acc
            dscale = (ctx.dscale * (g / hint)).to(ctx.scale_dtype)
            ctx.dscale = None
        nones = (None,) * 6
        wparam = holder.weight_param() if holder is not None else None
        tied = wparam is not None and holder.accumulate_into_grad and wparam.data_ptr() == ctx.w_ptr
        sync = getattr(holder, "sync", None) if tied else None
        if sync is not None:
            # data parallel: dX is rescaled here; the LM-head gradient is summed over ranks by the exchange -- already
            # running behind the last dX contraction when the multiplier was announced, started now otherwise
            outside = ctx.sync is None and sync.peer is not None
            if outside:
                # formed outside the exchange buffer (compute_cross_entropy): move it there before it is exchanged
                sync.before_dw_acc_write()  # the previous chunk's exchange may still read the buffer, here or remotely
                sync.dw_acc.copy_(dW_acc)
                dW_acc = sync.dw_acc
            sync.in_backward(dX, dW_acc, g, hint, formed_outside=outside)
            return (dX, None) + nones + (dscale, None, None)
        if tied:
            acc = wparam.grad is not None
            if not acc:
                wparam.grad = torch.empty_like(wparam)
            fold = getattr(holder, "fold", None)
            ps, fws = fold.dense_args(hint, dX.size(0)) if fold is not None else (None, None)
            ops.lmhead_ce_apply_grad(dX, dW_acc, g, hint, dW=wparam.grad, accumulate_dW=acc, post_scale=ps,
                                     fold_ws=fws)
            return (dX, None) + nones + (dscale, None, None)
        _, dW = ops.lmhead_ce_apply_grad(dX, dW_acc, g, hint)
        return (dX, dW) + nones + (dscale, None, None)


class _ChunkAlias(torch.autograd.Function):
    """The chunks a model hands over (``h.chunk(n, dim=1)``, tt/torchtune/modules/transformer.py:499-502) are adjacent
    views of one contiguous ``[b, s, d]`` tensor: forward returns that tensor again as an alias of the chunks' memory
    (no ``torch.cat`` copy), backward hands each chunk its slice of the gradient."""

    @staticmethod
    def forward(ctx, *chunks):
        c0 = chunks[0]
        ctx.sizes = [c.size(1) for c in chunks]
        total = sum(ctx.sizes)
        return torch.as_strided(c0.detach(), (c0.size(0), total, c0.size(2)), c0.stride(), c0.storage_offset())

    @staticmethod
    def backward(ctx, g):
        out, o = [], 0
        for n in ctx.sizes:
            out.append(g[:, o:o + n])
            o += n
        return tuple(out)


def _adjacent_chunks(chunks: Sequence[torch.Tensor]) -> bool:
    """True if the ``[b, s_i, d]`` chunks are consecutive slices along dim 1 of one contiguous ``[b, sum s_i, d]``."""
    c0 = chunks[0]
    if c0.dim() != 3:
        return False
    b, _, d = c0.shape
    total = sum(c.size(1) for c in chunks if c.dim() == 3)
    if c0.stride() != (total * d, d, 1):
        return False
    base = c0.untyped_storage().data_ptr()
    off = c0.storage_offset()
    for c in chunks:
        if (c.dim() != 3 or c.size(0) != b or c.size(2) != d or c.dtype != c0.dtype or c.device != c0.device
                or c.stride() != c0.stride() or c.untyped_storage().data_ptr() != base or c.storage_offset() != off):
            return False
        off += c.size(1) * d
    return True


@torch.compiler.disable
def fused_linear_cross_entropy(hidden: torch.Tensor, weight: torch.Tensor, labels: torch.Tensor,
                               ignore_index: int = -100, n_valid_total: torch.Tensor | None = None,
                               _holder=None, mode: str = "single_pass",
                               grad_hint: torch.Tensor | None = None, norm_scale: torch.Tensor | None = None,
                               norm_eps: float = 1e-6, reduction: str = "mean") -> torch.Tensor:
    """Functional form: ``hidden [N, D]`` bf16, ``weight [V, D]`` bf16, ``labels [N]`` int64 -> fp32 0-dim loss.

    ``mode``: ``"single_pass"`` (default; gradients formed during forward, 6 N V D flop) or ``"recompute"``
    (forward keeps only the per-row logsumexp, backward recomputes the logits tiles, 8 N V D flop, no stash).

    ``reduction``: ``"mean"`` over the valid rows (``n_valid_total``, an int64 device scalar, replaces the divisor:
    data parallelism) or ``"sum"`` (0 when every row is ignored; ce_chunked_output_loss.py:41-43).

    ``norm_scale`` (bf16 ``[D]`` Parameter) / ``norm_eps``: ``hidden`` is the PRE-norm state and the final RMSNorm
    (tt/torchtune/modules/rms_norm.py:38-43) runs inside the path (single-pass mode: fused into its row gather and
    dX scatter), with gradients for ``hidden`` and for the scale.

    The call is opaque to ``torch.compile`` (``torch.compiler.disable``): the kernels are reached through ctypes, so
    a compiled caller (``training.compile_loss`` wraps ``compute_cross_entropy``) runs it eagerly between graphs."""
    if reduction not in ("mean", "sum"):
        raise ValueError(f"ssi_b200: unknown reduction {reduction!r} (expected 'mean' or 'sum')")
    if mode == "recompute":
        if norm_scale is not None:  # no fused form in recompute mode: the stand-alone norm kernels in front
            from .model_hooks import _RMSNormFn

            hidden = _RMSNormFn.apply(hidden, norm_scale, norm_eps)
        sync = getattr(_holder, "sync", None)
        if sync is not None and reduction == "mean" and torch.is_grad_enabled():
            # data parallel: the divisor of the mean is the GLOBAL valid count (ssi_b200.sync, NCCL exchange)
            local = getattr(_holder, "local_counts", None)
            if local is None:
                local = ops.count_valid(labels, ignore_index)
            gcounts, _, ev = sync.global_counts(local)
            if ev is not None:
                torch.cuda.current_stream(hidden.device).wait_event(ev)
            n_valid_total = gcounts[1:2]
            _holder.last_global_counts = gcounts
        return _FusedLinearCE.apply(hidden, weight, labels, ignore_index, _holder, n_valid_total, reduction)
    if mode != "single_pass":
        raise ValueError(f"ssi_b200: unknown mode {mode!r} (expected 'single_pass' or 'recompute')")
    return _FusedLinearCESinglePass.apply(hidden, weight, labels, ignore_index, _holder, n_valid_total, grad_hint,
                                          torch.is_grad_enabled(), norm_scale, norm_eps, reduction)


class FusedLinearCrossEntropyLoss(nn.Module):
    """Same constructor attributes and call signature as ``CEWithChunkedOutputLoss``
    (tt/torchtune/modules/loss/ce_chunked_output_loss.py:30-33,45) plus the tied module whose ``.weight`` is
    the projection (tt/torchtune/modules/tied_linear.py:47-53).

    Args:
        tied_module: module owning the ``[V, D]`` bf16 weight (``model.tok_embeddings``).
        num_output_chunks: kept for ``model.set_num_output_chunks`` (ssi/trainer.py:303-304); the fused kernel
            does not need chunking, any value works.
        ignore_index: label value that contributes nothing (default -100).
        accumulate_into_grad: write dW directly into ``tied_module.weight.grad`` instead of returning it to
            autograd (saves a ``[V, D]`` temporary and an AccumulateGrad pass).
        mode: ``"single_pass"`` (default) or ``"recompute"``, see :func:`fused_linear_cross_entropy`;
            overridable with the environment variable ``SSI_B200_MODE``.

    Attribute ``grad_hint`` (fp32 device scalar or None = 1.0): the multiplier the caller will apply to the
    loss before ``backward()`` (ssi/trainer.py:393 uses ``num_tokens_iter``). Purely an optimisation: with the
    right hint single-pass gradients are rounded to bf16 once; any other upstream gradient is still exact up to
    one extra bf16 rounding of dX.

    Data parallelism (``install(model, process_group=...)`` sets ``sync``, see :mod:`ssi_b200.sync`): the training
    forward returns ``local loss sum / GLOBAL valid count`` (the rank's share of the global mean) and the tied gradient
    is summed over ranks.  ``local_counts`` (int64 ``[2]`` device tensor {unshifted, shifted} or None) lets the trainer
    hand over counts it already has; ``hint_is_global_count = True`` announces the multiplier "global number of
    unshifted valid tokens" (ssi/trainer.py:391-393 on the global batch), which the counts exchange produces on the
    device; ``last_global_counts`` holds the summed counts of the last training forward.  Forward-only calls (eval,
    ssi/eval.py:27-33) never use any of this: every rank gets the mean over ITS batch.
    """

    def __init__(self, tied_module: nn.Module, num_output_chunks: int = 8, ignore_index: int = -100,
                 accumulate_into_grad: bool = True, mode: str | None = None):
        super().__init__()
        if not hasattr(tied_module, "weight"):
            raise AttributeError(
                "Provided module does not have attribute 'weight'. Please check your tied_module."
            )
        # not registered as a submodule: the weight must stay owned by the model (state-dict keys unchanged)
        object.__setattr__(self, "_tied", tied_module)
        self.num_output_chunks = num_output_chunks
        self.ignore_index = ignore_index
        self.accumulate_into_grad = accumulate_into_grad
        self.mode = mode or os.environ.get("SSI_B200_MODE", "single_pass")
        if self.mode not in ("single_pass", "recompute"):
            raise ValueError(f"ssi_b200: unknown mode {self.mode!r} (expected 'single_pass' or 'recompute')")
        self.grad_hint: torch.Tensor | None = None
        self.last_loss_sum: torch.Tensor | None = None
        self.last_n_valid: torch.Tensor | None = None
        self.last_global_counts: torch.Tensor | None = None
        object.__setattr__(self, "sync", None)  # ssi_b200.sync.TiedGradSync (data parallelism); not a submodule
        object.__setattr__(self, "fold", None)  # ssi_b200.fold.TiedGradFold (install(fold_scale_grads=True))
        self.local_counts: torch.Tensor | None = None
        self.hint_is_global_count: bool = False
        object.__setattr__(self, "_norm", None)  # fused final RMSNorm (set_fused_norm); not a registered submodule

    @property
    def tied_module(self) -> nn.Module:
        return self._tied

    @property
    def grad_hint_mismatch(self):
        return self.sync.grad_hint_mismatch if self.sync is not None else None

    def set_fused_norm(self, norm: nn.Module | None) -> None:
        """Take over the model's final RMSNorm (``model.norm``, applied at tt/torchtune/modules/transformer.py:653):
        ``norm`` must expose ``scale`` (bf16 ``[D]`` Parameter) and ``eps``.  The model then hands PRE-norm hidden
        states to the loss (``install(model, fuse_norm=True)`` swaps ``model.norm`` for a passthrough that still
        owns the Parameter, so ``norm.scale`` stays in the state dict and the optimizer)."""
        if norm is not None and not (hasattr(norm, "scale") and hasattr(norm, "eps")):
            raise AttributeError("ssi_b200: the fused norm needs `scale` and `eps` attributes (torchtune RMSNorm)")
        object.__setattr__(self, "_norm", norm)

    def _norm_args(self):
        n = self._norm
        return (None, 1e-6) if n is None else (n.scale, float(n.eps))

    def weight_param(self):
        w = self._tied.weight
        return w if isinstance(w, nn.Parameter) else None

    def compute_cross_entropy(self, hidden: torch.Tensor, labels: torch.Tensor, normalize: bool = True):
        """SUM-reduced CE of one chunk (ce_chunked_output_loss.py:35-43: ``F.cross_entropy(logits.float(), labels,
        ignore_index, reduction="sum")``; 0 -- not NaN -- when every label of the chunk is ignored).  The reference
        passes the chunk's logits; with the passthrough output installed the model hands over HIDDEN states and the
        projection happens here.  ``training.compile_loss`` wraps exactly this method
        (tt/torchtune/training/_compile.py:74-77); the fused call inside is opaque to ``torch.compile``."""
        return fused_linear_cross_entropy(hidden.reshape(-1, hidden.size(-1)), self._tied.weight,
                                          labels.reshape(-1).contiguous(), self.ignore_index, None, self, self.mode,
                                          None, *self._norm_args(), reduction="sum")

    def forward(self, hidden: Union[torch.Tensor, Sequence[torch.Tensor]], labels: torch.Tensor) -> torch.Tensor:
        W = self._tied.weight
        if isinstance(hidden, (list, tuple)):
            # list of [b, s_c, D] chunks along dim 1 (tt/torchtune/modules/transformer.py:499-502)
            if len(hidden) == 1:
                h = hidden[0]
            elif _adjacent_chunks(hidden):
                h = _ChunkAlias.apply(*hidden)  # adjacent views of one contiguous tensor: no copy
            else:
                h = torch.cat(list(hidden), dim=1)
        else:
            h = hidden
        D = h.size(-1)
        X = h.reshape(-1, D)
        if not X.is_contiguous():
            X = X.contiguous()
        lab = labels.reshape(-1)
        if not lab.is_contiguous():
            lab = lab.contiguous()
        if lab.numel() != X.size(0):
            raise ValueError(f"labels ({tuple(labels.shape)}) do not match hidden rows ({X.size(0)})")
        return fused_linear_cross_entropy(X, W, lab, self.ignore_index, None, self, self.mode, self.grad_hint,
                                          *self._norm_args())


def make_torchtune_subclass():
    """If torchtune is importable, return a subclass of BOTH CEWithChunkedOutputLoss and the fused loss so that
    ``isinstance(loss_fn, CEWithChunkedOutputLoss)`` in ssi/trainer.py:303 keeps holding; else None."""
    try:  # torchtune is not importable in the build image: tests/test_host_logic.py stubs the module path
        from torchtune.modules.loss import CEWithChunkedOutputLoss
    except Exception:
        return None

    class FusedCEWithChunkedOutputLoss(FusedLinearCrossEntropyLoss, CEWithChunkedOutputLoss):
        pass

    return FusedCEWithChunkedOutputLoss
