"""One micro-batch of the graded path, driven the way ``Trainer._train_step`` drives the reference
(ssi/trainer.py:385-395) THROUGH THE DROP-IN API (``install()``), with the 16 decoder layers cut out:

    h0 = model.tok_embeddings(tokens)                          # a6 forward gather (+ the id sort its backward needs)
    num_tokens_iter = (labels != ignore).sum()                 # ssi/trainer.py:391 (unshifted labels)
    labels' = shift(labels)                                    # ssi/loss.py:16
    loss = loss_fn(hidden, labels')                            # fused projection + CE  (K1 .. dW, dX)
    (loss * num_tokens_iter).backward()                        # dX; LM-head part of W.grad
    [the decoder backward would run here: it consumes dX and produces dH0]
    h0.backward(dH0)                                           # K3: W.grad[tokens] += dH0  -- AFTER the LM head

``hidden`` stands for the final hidden state and ``dH0`` for the gradient arriving at the embedding output; both are
inputs here because the decoder layers are outside the path.  The order is causal: the embedding gradient is handed
over only after dX was returned, exactly as in a real step.  With a process group, ``install(process_group=...)``
wires the exchange of the tied gradient (:mod:`ssi_b200.sync`): counts at the start, the LM-head part behind the last
dX contraction, the embedding part after the embedding backward.
"""
from __future__ import annotations

import torch
from torch import nn

from . import ops
from .model_hooks import install


class _PathModel(nn.Module):
    """What is left of the reference's TransformerDecoder on the path: the tied embedding and the output slot."""

    def __init__(self, weight: nn.Parameter):
        super().__init__()
        self.tok_embeddings = nn.Embedding(weight.size(0), weight.size(1), device="meta", dtype=weight.dtype)
        self.tok_embeddings.weight = weight
        self.output = None


class PathStep:
    """``exchange``: how the tied gradient is summed over ranks when a process group is given — ``"peer"`` (default
    for single-pass mode: the library's own NVLink peer-memory kernels) or ``"nccl"`` (bf16 ``all_reduce`` of the
    micro-batch's contribution after the embedding backward).  Environment override: ``SSI_B200_EXCHANGE``.
    ``fold_scale_grads``: every call is one optimizer step; ``weight.grad`` then holds the gradient ALREADY scaled by
    ``1 / num_tokens_step`` (ssi/trainer.py:404) and ``self.fold.sqnorm()`` its sum of squares (:mod:`ssi_b200.fold`)."""

    def __init__(self, weight: nn.Parameter | torch.Tensor, ignore_index: int = -100, num_output_chunks: int = 8,
                 process_group=None, mode: str | None = None, exchange: str | None = None,
                 max_tokens_per_rank: int = 65536, fold_scale_grads: bool = False):
        if not isinstance(weight, nn.Parameter):
            weight = nn.Parameter(weight)
        self.model = _PathModel(weight)
        self.world = 1
        if process_group is not None:
            import torch.distributed as dist

            self.world = dist.get_world_size(process_group)
        use_pg = process_group if (process_group is not None and (self.world > 1 or exchange in ("peer", "nvls"))) else None
        self.loss_fn = install(self.model, num_output_chunks, ignore_index, mode=mode, process_group=use_pg,
                               exchange=exchange, max_tokens_per_rank=max_tokens_per_rank,
                               fold_scale_grads=fold_scale_grads)
        # one micro-batch per optimizer step: scale_grads(model, 1 / num_tokens_step) (ssi/trainer.py:404) is the
        # reciprocal of the announced multiplier and rides in the kernels that write the tied gradient
        self.fold = self.loss_fn.fold
        self.fold_every_step = self.fold is not None  # False: the caller arms the fold itself (accumulation windows)
        self.emb = self.model.tok_embeddings
        self.sync = self.loss_fn.sync
        self.exchange = self.sync.mode if self.sync is not None else None
        self.ignore_index = ignore_index
        self.pg = process_group

    def set_fold(self, enabled: bool) -> None:
        """Switch the scale_grads / grad-norm fold on or off for the following steps."""
        from .model_hooks import attach_fold

        self.fold = attach_fold(self.loss_fn, self.emb, enabled)
        self.fold_every_step = self.fold is not None

    @property
    def weight(self) -> nn.Parameter:
        return self.emb.weight

    def zero_grad(self, set_to_none: bool = True) -> None:
        if set_to_none:
            self.weight.grad = None
        elif self.weight.grad is not None:
            self.weight.grad.zero_()

    def forward_only(self, hidden: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        """Eval path (ssi/eval.py:26-33): loss of THIS rank's batch under inference_mode, no backward, no exchange."""
        with torch.inference_mode():
            shifted, _ = ops.label_shift_count(labels, self.ignore_index)
            return self.loss_fn(hidden, shifted)

    def __call__(self, tokens: torch.Tensor, labels: torch.Tensor, hidden: torch.Tensor, dH0: torch.Tensor):
        """Returns ``(loss_batch, dX, counts)``: loss_batch = loss * num_tokens_iter (fp32 0-dim, what
        ssi/trainer.py:393-395 accumulates; with a process group: this rank's share, global token numbers), dX bf16
        like ``hidden``, counts int64 [2] (unshifted, shifted; LOCAL).  ``self.weight.grad`` holds LM-head +
        embedding gradient (summed over ranks when a process group is set)."""
        loss_fn = self.loss_fn
        if self.sync is not None:
            self.sync.check()
        if self.fold_every_step:
            self.fold.arm()  # post_scale = 1 / num_tokens_iter (global with a process group)
        h0 = self.emb(tokens)                                   # transformer.py:628
        shifted, counts = ops.label_shift_count(labels, self.ignore_index)
        hidden = hidden.detach().requires_grad_(True)
        if self.sync is not None:
            # every rank scales by the GLOBAL token numbers, so the gradients are those of the single-process loss on
            # the concatenated batch; the counts travel inside the loss forward (no collective in front of it)
            loss_fn.local_counts = counts
            loss_fn.hint_is_global_count = True
            loss = loss_fn(hidden, shifted)
            n_iter = loss_fn.last_global_counts[0].to(torch.float32)
        else:
            n_iter = counts[0].to(torch.float32)
            # announce the multiplier of ssi/trainer.py:393 so single-pass mode rounds its gradients once
            loss_fn.grad_hint = n_iter.reshape(1)
            loss = loss_fn(hidden, shifted)
        loss_batch = loss * n_iter
        loss_batch.backward()                                   # dX (and, single GPU, the LM-head part of .grad)
        dX = hidden.grad
        # ---- the decoder backward would run here: dX in, dH0 out
        h0.backward(dH0.reshape(h0.shape))                      # tied-embedding backward: the end of the micro-batch
        return loss_batch.detach(), dX, counts

    def close(self) -> None:
        """Release the peer exchange buffers (collective: every rank must call it)."""
        if self.sync is not None:
            self.sync.check(confirm_multiplier=True)
            self.sync.close()
