"""``scale_grads`` and the tied weight's share of ``clip_grad_norm_`` folded into the path.

The reference ends an accumulation window with two more passes over every gradient
(ssi/trainer.py:404-408: ``scale_grads(model, 1 / num_tokens_step)``, tt/torchtune/training/_grad_scaler.py:25-31, then
``clip_grad_norm_``).  For the tied weight -- 534 MB at V = 133,258 -- the path already rewrites every element once
(the dense site: ``apply_dw`` or the peer reduce) and every touched row once more (the sparse site: the embedding
backward), so both passes ride along: the scaler is applied in fp32 before the one rounding to bf16 and each kernel
leaves its share of sum(g^2).  :class:`TiedGradFold` is created by ``install(..., fold_scale_grads=True)``; a trainer

    loss_fn.fold.arm(torch.tensor(1 / num_tokens_step))      # before the forward of the window's LAST micro-batch
    ... forward / backward of that micro-batch ...
    ssi_b200.fold.scale_grads(model, scaler, loss_fn)          # skips the tied weight (already scaled)
    total = ssi_b200.fold.grad_norm(model, loss_fn)            # tied weight: the folded sum of squares

``arm()`` without a scaler means "the reciprocal of the announced loss multiplier" (``loss_fn.grad_hint``; with one
micro-batch per step that IS 1 / num_tokens_step, ssi/trainer.py:391-393), which never leaves the device.
"""
from __future__ import annotations

import torch
from torch import nn

from . import ops


class TiedGradFold:
    def __init__(self, weight: nn.Parameter):
        self.weight = weight
        self.device = weight.device
        self.post_scale = torch.ones(1, dtype=torch.float32, device=self.device)
        self.ws: torch.Tensor | None = None
        self._rows = 0
        self.armed = False      # the next micro-batch folds
        self.folded = False     # the last micro-batch folded: .grad is scaled, sqnorm() is valid
        self._from_hint = False
        self._dense_done = False

    def arm(self, scaler: torch.Tensor | float | None = None) -> None:
        """Fold the scaling (and the sum of squares) into the NEXT micro-batch; call before its forward."""
        self.armed, self.folded, self._dense_done = True, False, False
        self._from_hint = scaler is None
        if scaler is not None:
            s = torch.as_tensor(scaler, dtype=torch.float32).reshape(1)
            self.post_scale.copy_(s, non_blocking=True)

    def disarm(self) -> None:
        self.armed = False

    def ensure(self, n_rows: int) -> None:
        """Size the scratch for micro-batches of ``n_rows`` token rows (call on the stream that owns allocations)."""
        if self.armed and (self.ws is None or n_rows > self._rows):
            self.ws = ops.grad_fold_workspace(n_rows, self.weight.size(1), self.device)
            self._rows = n_rows

    def dense_args(self, hint: torch.Tensor, n_rows: int):
        """``(post_scale, fold_ws)`` for the dense site of an armed micro-batch, else ``(None, None)``.  Enqueues on
        the current stream (which must be ordered after ``hint``)."""
        if not self.armed:
            return None, None
        self.ensure(n_rows)
        if self._from_hint:
            h = hint.reshape(1).to(torch.float32)
            self.post_scale.copy_(torch.where(h > 0, 1.0 / h, torch.zeros_like(h)))
        self._dense_done = True
        return self.post_scale, self.ws

    def sparse_args(self, n_rows: int):
        """``(post_scale, fold_ws)`` for the sparse site (the embedding backward ends the micro-batch)."""
        if not self.armed:
            return None, None
        if not self._dense_done:
            raise RuntimeError("ssi_b200.fold: the embedding backward of an armed micro-batch ran before the loss "
                               "backward (the dense site starts a fold)")
        if n_rows > self._rows:
            raise RuntimeError(f"ssi_b200.fold: the embedding saw {n_rows} token rows, the loss only {self._rows}")
        self.armed, self.folded = False, True
        return self.post_scale, self.ws

    def sqnorm(self) -> torch.Tensor:
        """fp32 ``[1]`` device tensor: sum of squares of the tied gradient after the folded micro-batch."""
        if not self.folded:
            raise RuntimeError("ssi_b200.fold: no folded micro-batch (arm() before its forward)")
        return ops.grad_fold_sqnorm(self.ws)


def _fold_of(loss_fn) -> TiedGradFold | None:
    f = getattr(loss_fn, "fold", None) if loss_fn is not None else None
    return f if (f is not None and f.folded) else None


def scale_grads(model: nn.Module, scaler: torch.Tensor, loss_fn=None) -> None:
    """Mirror of tt/torchtune/training/_grad_scaler.py:12-31 (``p.grad *= scaler`` for every parameter); the tied
    weight is skipped when its last micro-batch folded the scaling."""
    f = _fold_of(loss_fn)
    device = None
    for p in model.parameters():
        if not device:
            device = p.device
            scaler = scaler.to(device)
        if p.grad is not None and not (f is not None and p is f.weight):
            p.grad *= scaler


def grad_norm(model: nn.Module, loss_fn=None) -> torch.Tensor:
    """Total L2 norm of the gradients (what ``clip_grad_norm_`` computes first, ssi/trainer.py:405-408); the tied
    weight contributes its folded sum of squares instead of one more pass."""
    f = _fold_of(loss_fn)
    sq = None
    for p in model.parameters():
        if p.grad is None:
            continue
        t = f.sqnorm()[0] if (f is not None and p is f.weight) else p.grad.float().pow(2).sum()
        sq = t if sq is None else sq + t.to(sq.device)
    return torch.sqrt(sq) if sq is not None else torch.zeros(())
