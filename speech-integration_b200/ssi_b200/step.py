"""One micro-batch of the graded path, driven the way ``Trainer._train_step`` drives the reference
(ssi/trainer.py:385-395) with the 16 decoder layers cut out:

    num_tokens_iter = (labels != ignore).sum()                 # ssi/trainer.py:391 (unshifted labels)
    labels' = shift(labels)                                    # ssi/loss.py:16
    loss = loss_fn(hidden, labels')                            # fused projection + CE  (K1)
    (loss * num_tokens_iter).backward()                        # K2: dX, dW into W.grad
    tok_embeddings backward: W.grad[tokens] += dH0             # K3
    [data parallel] all-reduce W.grad, loss_sum, n_valid       # K4 (ssi_b200.dist)

``hidden`` stands for the final normed hidden state and ``dH0`` for the gradient arriving at the embedding
output; both are inputs here because the decoder layers are outside the path.
"""
from __future__ import annotations

import torch
from torch import nn

from . import ops
from .loss import FusedLinearCrossEntropyLoss


class PathStep:
    """``exchange``: how the tied gradient is summed over ranks when a process group is given —
    ``"peer"`` (default for single-pass mode): the library's own NVLink peer-memory reduce (ssi_b200.peer), started
    on a side stream as soon as the last dW contraction is done, so it runs behind the last dX contraction;
    the sum over ranks is taken in fp32 and rounded once.  ``"nccl"``: bf16 ``all_reduce`` after the step (the
    plain baseline).  Environment override: ``SSI_B200_EXCHANGE``."""

    def __init__(self, weight: nn.Parameter | torch.Tensor, ignore_index: int = -100, num_output_chunks: int = 8,
                 process_group=None, mode: str | None = None, exchange: str | None = None):
        if not isinstance(weight, nn.Parameter):
            weight = nn.Parameter(weight)
        self.emb = nn.Embedding(weight.size(0), weight.size(1), device="meta", dtype=weight.dtype)
        self.emb.weight = weight
        self.loss_fn = FusedLinearCrossEntropyLoss(self.emb, num_output_chunks, ignore_index, mode=mode)
        self.ignore_index = ignore_index
        self.pg = process_group
        self.world = 1
        if process_group is not None:
            import torch.distributed as dist

            self.world = dist.get_world_size(process_group)
        self.exchange = None
        self.peer = None
        if process_group is not None and (self.world > 1 or exchange == "peer"):
            import os

            self.exchange = exchange or os.environ.get("SSI_B200_EXCHANGE") or (
                "peer" if self.loss_fn.mode == "single_pass" else "nccl")
            if self.exchange not in ("peer", "nccl"):
                raise ValueError(f"ssi_b200: unknown exchange {self.exchange!r} (expected 'peer' or 'nccl')")
            if self.exchange == "peer":
                if self.loss_fn.mode != "single_pass":
                    raise ValueError("ssi_b200: exchange='peer' needs mode='single_pass'")
                from .peer import TiedGradExchange

                dev = weight.device
                try:
                    self.peer = TiedGradExchange(weight.size(0), weight.size(1), process_group, dev)
                except RuntimeError as e:
                    if exchange == "peer" or os.environ.get("SSI_B200_EXCHANGE") == "peer":
                        raise  # asked for explicitly
                    import warnings

                    # the failure was agreed on by all ranks (TiedGradExchange), so all of them land here together
                    warnings.warn(f"{e}; using the NCCL all-reduce of the tied gradient instead")
                    self.exchange = "nccl"
            if self.peer is not None:
                # SSI_B200_PEER_SERIAL=1 (diagnostics): run the exchange on the main stream after the step instead
                self.side = (torch.cuda.current_stream(dev) if os.environ.get("SSI_B200_PEER_SERIAL") == "1"
                             else torch.cuda.Stream(dev))
                self.dw_ready, self.dw_init, self.sorted = (torch.cuda.Event() for _ in range(3))
                for e in (self.dw_ready, self.dw_init, self.sorted):
                    e.record(torch.cuda.current_stream(dev))  # creates the CUDA handle
                self.loss_fn.dw_acc_buffer = self.peer.dw_acc
                self.loss_fn.dw_ready_event = self.dw_ready
                self.loss_fn.dw_acc_initialised = True
                self.loss_fn.dw_acc_init_event = self.dw_init
                self.loss_fn.defer_dw = True

    @property
    def weight(self) -> nn.Parameter:
        return self.emb.weight

    def zero_grad(self, set_to_none: bool = True) -> None:
        if set_to_none:
            self.weight.grad = None
        elif self.weight.grad is not None:
            self.weight.grad.zero_()

    def forward_only(self, hidden: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        """Eval path (ssi/eval.py:26-33): loss under inference_mode, no backward."""
        with torch.inference_mode():
            shifted, _ = ops.label_shift_count(labels, self.ignore_index)
            return self.loss_fn(hidden, shifted)

    def __call__(self, tokens: torch.Tensor, labels: torch.Tensor, hidden: torch.Tensor, dH0: torch.Tensor):
        """Returns ``(loss_batch, dX, counts)``: loss_batch = mean loss * num_tokens_iter (fp32 0-dim, what
        ssi/trainer.py:393-395 accumulates), dX bf16 like ``hidden``, counts int64 [2] (unshifted, shifted).
        ``self.weight.grad`` holds LM-head + embedding gradient (all-reduced when a process group is set)."""
        shifted, counts = ops.label_shift_count(labels, self.ignore_index)
        hidden = hidden.detach().requires_grad_(True)
        if self.peer is not None:
            # Embedding gradient FIRST: sort the ids here, then on the side stream (beside the first logits
            # contraction) dW_acc = 0 + scatter-add of dH0; every dW contraction then adds to it (the first one
            # waits for `dw_init`), and the exchange is the only thing left for the tail of the step.
            main = torch.cuda.current_stream(hidden.device)
            D_ = dH0.size(-1)
            sort_ws = ops.embedding_grad_sort(tokens.reshape(-1), self.weight.size(0), self.weight.size(1))
            self.sorted.record(main)
            self.side.wait_event(self.sorted)
            with torch.cuda.stream(self.side):
                self.peer.dw_acc.zero_()
                ops.embedding_grad_scatter_sorted_f32(dH0.reshape(-1, D_), sort_ws, self.peer.dw_acc)
                self.dw_init.record(self.side)
        gcounts = counts
        if self.world > 1:
            from . import dist as ddist

            # global counts: every rank scales by the GLOBAL token numbers, so gradients are those of the
            # single-process loss on the concatenated batch
            gcounts = ddist.global_counts(counts, self.pg)
            self.loss_fn.n_valid_total = gcounts[1:2]
        n_iter = gcounts[0].to(torch.float32)
        # announce the multiplier of ssi/trainer.py:393 so single-pass mode rounds its gradients once
        self.loss_fn.grad_hint = n_iter.reshape(1)
        loss = self.loss_fn(hidden, shifted)
        loss_batch = loss * n_iter
        D = dH0.size(-1)
        if self.peer is not None:
            # The forward above recorded `dw_ready` right after the last dW contraction.  Side stream, behind the
            # last dX contraction and the rest of the main stream: the peer reduce of the fp32 sums (LM head +
            # embedding) over ranks, one rounding, straight into .grad.
            accumulate = self.weight.grad is not None
            if not accumulate:
                self.weight.grad = torch.empty_like(self.weight)
            hint = self.loss_fn.grad_hint
            self.side.wait_event(self.dw_ready)
            with torch.cuda.stream(self.side):
                self.peer.exchange(self.weight.grad, hint, hint, accumulate=accumulate)
            loss_batch.backward()  # dX rescale only (a device-side no-op: the multiplier was announced)
            main.wait_stream(self.side)
            return loss_batch.detach(), hidden.grad, counts
        loss_batch.backward()
        ops.embedding_grad_scatter(dH0.reshape(-1, D), tokens.reshape(-1), self.weight.grad)
        if self.world > 1:
            from . import dist as ddist

            ddist.allreduce_tied_grad(self.weight.grad, self.pg)
        return loss_batch.detach(), hidden.grad, counts

    def close(self) -> None:
        """Release the peer exchange buffers (collective: every rank must call it)."""
        if self.peer is not None:
            self.loss_fn.dw_acc_buffer = None
            self.peer.close()
            self.peer = None
