"""Drop-in replacement for torchtune's ``CEWithChunkedOutputLoss`` that fuses the tied output projection.

Reference call signature (ssi/trainer.py:299-304, ssi/loss.py:7-22): the trainer builds the loss, reads
``num_output_chunks`` / ``ignore_index`` and calls ``loss_fn(model_output, labels)`` where ``model_output`` is a
list of ``num_output_chunks`` tensors ``[b, s/n, *]`` (labels ``[b, s]``) or a single 2-D tensor (labels
``[b*s]``).  With :class:`ssi_b200.HiddenPassthroughOutput` installed as ``model.output`` the model returns
the final normed HIDDEN states in those positions and this module does projection + CE in one fused pass.
"""
from __future__ import annotations

import os
from typing import List, Sequence, Union

import torch
from torch import nn

from . import ops


class _FusedLinearCE(torch.autograd.Function):
    """loss = mean_{valid rows} CE(X W^T, labels). Saves only X, W, labels and the per-row logsumexp."""

    @staticmethod
    def forward(ctx, X, W, labels, ignore_index, holder, n_valid_override):
        loss_out, lse, n_valid = ops.lmhead_ce_fwd(X, W, labels, ignore_index)
        ctx.ignore_index = ignore_index
        ctx.holder = holder
        n_div = n_valid if n_valid_override is None else n_valid_override
        ctx.save_for_backward(X, W, labels, lse, n_div)
        if holder is not None:
            holder.last_loss_sum = loss_out[0:1]
            holder.last_n_valid = n_valid
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
            return dX, None, None, None, None, None
        dX, dW = ops.lmhead_ce_bwd(X, W, labels, lse, g, n_div, ctx.ignore_index)
        return dX, dW, None, None, None, None


_ONES: dict = {}


def _one(device: torch.device) -> torch.Tensor:
    t = _ONES.get(device)
    if t is None:
        t = torch.ones(1, dtype=torch.float32, device=device)
        _ONES[device] = t
    return t


class _FusedLinearCESinglePass(torch.autograd.Function):
    """Single-pass mode: when a gradient is needed, forward also forms dX and the fp32 dW sum (each logits tile
    is computed once; a bounded bf16 block of logits/dlogits is the only scratch). Backward only rescales by
    ``grad_out / grad_hint`` (1 when the trainer announced its multiplier) and rounds dW once into ``.grad``."""

    @staticmethod
    def forward(ctx, X, W, labels, ignore_index, holder, n_valid_override, grad_hint, grad_enabled=True,
                norm_scale=None, norm_eps=1e-6):
        ctx.holder = holder
        ctx.dscale = None
        # `grad_enabled` is sampled by the caller: inside Function.forward grad mode is always off, and under
        # torch.no_grad() / inference_mode() (the eval path, ssi/eval.py:27-33) needs_input_grad still reports the
        # Parameter -- forming gradients there would triple the work of a forward-only call
        ctx.with_grad = bool(grad_enabled and (ctx.needs_input_grad[0] or ctx.needs_input_grad[1]
                                               or (norm_scale is not None and ctx.needs_input_grad[8])))
        if ctx.with_grad:
            hint = _one(X.device) if grad_hint is None else grad_hint.detach().to(torch.float32).reshape(1)
            out = ops.lmhead_ce_fwd_grad(
                X, W, labels, hint, n_valid_override, ignore_index,
                dW_acc=getattr(holder, "dw_acc_buffer", None), dw_ready_event=getattr(holder, "dw_ready_event", None),
                accumulate_dW_acc=getattr(holder, "dw_acc_initialised", False),
                dw_acc_init_event=getattr(holder, "dw_acc_init_event", None),
                norm_scale=norm_scale.detach() if norm_scale is not None else None, norm_eps=norm_eps)
            loss_out, lse, n_valid, dX, dW_acc = out[:5]
            if norm_scale is not None:  # X was the PRE-norm state: dX is its gradient, dscale the scale's
                ctx.dscale, ctx.scale_dtype = out[5], norm_scale.dtype
            ctx.hint, ctx.dX, ctx.dW_acc, ctx.w_ptr = hint, dX, dW_acc, W.data_ptr()
        else:
            if norm_scale is not None:  # forward only (eval): the stand-alone norm kernel, then K1
                X = ops.rmsnorm_fwd(X, norm_scale.detach(), norm_eps, want_rstd=False)[0]
            loss_out, lse, n_valid = ops.lmhead_ce_fwd(X, W, labels, ignore_index)
        if holder is not None:
            holder.last_loss_sum = loss_out[0:1]
            holder.last_n_valid = n_valid
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
        wparam = holder.weight_param() if holder is not None else None
        if holder is not None and getattr(holder, "defer_dw", False):
            # data-parallel peer exchange (ssi_b200.peer / PathStep): the SUMMED fp32 dW is rounded into .grad by
            # the exchange, which already runs behind the last dX contraction with the announced multiplier;
            # here only dX is rescaled, and a wrong announcement is left visible instead of silently ignored.
            ops.lmhead_ce_apply_grad(dX, dW_acc, g, hint, skip_dW=True)
            holder.grad_hint_mismatch = (g != hint)
            return dX, None, None, None, None, None, None, None, dscale, None
        if wparam is not None and holder.accumulate_into_grad and wparam.data_ptr() == ctx.w_ptr:
            acc = wparam.grad is not None
            if not acc:
                wparam.grad = torch.empty_like(wparam)
            ops.lmhead_ce_apply_grad(dX, dW_acc, g, hint, dW=wparam.grad, accumulate_dW=acc)
            return dX, None, None, None, None, None, None, None, dscale, None
        _, dW = ops.lmhead_ce_apply_grad(dX, dW_acc, g, hint)
        return dX, dW, None, None, None, None, None, None, dscale, None


def fused_linear_cross_entropy(hidden: torch.Tensor, weight: torch.Tensor, labels: torch.Tensor,
                               ignore_index: int = -100, n_valid_total: torch.Tensor | None = None,
                               _holder=None, mode: str = "single_pass",
                               grad_hint: torch.Tensor | None = None, norm_scale: torch.Tensor | None = None,
                               norm_eps: float = 1e-6) -> torch.Tensor:
    """Functional form: ``hidden [N, D]`` bf16, ``weight [V, D]`` bf16, ``labels [N]`` int64 -> fp32 0-dim loss.

    ``mode``: ``"single_pass"`` (default; gradients formed during forward, 6 N V D flop) or ``"recompute"``
    (forward keeps only the per-row logsumexp, backward recomputes the logits tiles, 8 N V D flop, no stash).

    ``norm_scale`` (bf16 ``[D]`` Parameter) / ``norm_eps``: ``hidden`` is the PRE-norm state and the final RMSNorm
    (tt/torchtune/modules/rms_norm.py:38-43) runs inside the path (single-pass mode: fused into its row gather and
    dX scatter), with gradients for ``hidden`` and for the scale."""
    if mode == "recompute":
        if norm_scale is not None:  # no fused form in recompute mode: the stand-alone norm kernels in front
            from .model_hooks import _RMSNormFn

            hidden = _RMSNormFn.apply(hidden, norm_scale, norm_eps)
        return _FusedLinearCE.apply(hidden, weight, labels, ignore_index, _holder, n_valid_total)
    if mode != "single_pass":
        raise ValueError(f"ssi_b200: unknown mode {mode!r} (expected 'single_pass' or 'recompute')")
    return _FusedLinearCESinglePass.apply(hidden, weight, labels, ignore_index, _holder, n_valid_total, grad_hint,
                                          torch.is_grad_enabled(), norm_scale, norm_eps)


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
        self.n_valid_total: torch.Tensor | None = None  # set by ssi_b200.dist for data parallelism
        self.last_loss_sum: torch.Tensor | None = None
        self.last_n_valid: torch.Tensor | None = None
        # data-parallel peer exchange hooks (set by ssi_b200.step.PathStep): the fp32 accumulator to write into,
        # the event to record once it is final, and "leave dW to the exchange" for backward
        self.dw_acc_buffer: torch.Tensor | None = None
        self.dw_ready_event: torch.cuda.Event | None = None
        self.dw_acc_initialised: bool = False  # the caller pre-fills dw_acc_buffer (embedding gradient) every step
        self.dw_acc_init_event: torch.cuda.Event | None = None
        self.defer_dw: bool = False
        self.grad_hint_mismatch: torch.Tensor | None = None
        object.__setattr__(self, "_norm", None)  # fused final RMSNorm (set_fused_norm); not a registered submodule

    @property
    def tied_module(self) -> nn.Module:
        return self._tied

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
        """Sum-reduced CE of one chunk of hidden states (mirrors ce_chunked_output_loss.py:35-43, which takes
        logits): returned so that ``training.compile_loss``-style wrappers that call it keep working."""
        loss = fused_linear_cross_entropy(hidden.reshape(-1, hidden.size(-1)), self._tied.weight,
                                          labels.reshape(-1), self.ignore_index, None, self, self.mode, self.grad_hint,
                                          *self._norm_args())
        return loss * self.last_n_valid[0].to(torch.float32)

    def forward(self, hidden: Union[torch.Tensor, Sequence[torch.Tensor]], labels: torch.Tensor) -> torch.Tensor:
        W = self._tied.weight
        if isinstance(hidden, (list, tuple)):
            # list of [b, s_c, D] chunks along dim 1 (tt/torchtune/modules/transformer.py:499-502)
            h = hidden[0] if len(hidden) == 1 else torch.cat(list(hidden), dim=1)
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
        return fused_linear_cross_entropy(X, W, lab, self.ignore_index, self.n_valid_total, self, self.mode,
                                          self.grad_hint, *self._norm_args())


def make_torchtune_subclass():
    """If torchtune is importable, return a subclass of BOTH CEWithChunkedOutputLoss and the fused loss so that
    ``isinstance(loss_fn, CEWithChunkedOutputLoss)`` in ssi/trainer.py:303 keeps holding; else None."""
    try:  # pragma: no cover - torchtune is not importable in the build image
        from torchtune.modules.loss import CEWithChunkedOutputLoss
    except Exception:
        return None

    class FusedCEWithChunkedOutputLoss(FusedLinearCrossEntropyLoss, CEWithChunkedOutputLoss):  # pragma: no cover
        pass

    return FusedCEWithChunkedOutputLoss
