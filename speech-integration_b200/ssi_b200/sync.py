"""Data parallelism of the path behind the drop-in API, in the order a training step produces the pieces.

The reference trains on one GPU and only scaffolds DP at the data level (``DistributedSampler``,
ssi/data/__init__.py:82-83); the semantics to match are those of torchtune's distributed recipe
(tt/recipes/full_finetune_distributed.py:844-851,867): token counts and losses are summed over ranks and the
gradients are those of the global mean.  :class:`TiedGradSync` is attached to the fused loss and to the tied embedding
by ``install(model, process_group=...)`` and drives, per micro-batch:

1. **counts** — at the start of the loss forward the local token counts are summed over ranks (peer mode: 8-byte
   peer stores on a side stream beside the first logits contraction, no collective call on the critical path; the
   main stream only waits for them right before the first softmax-grad pass, their first consumer);
2. **LM-head part** — single-pass mode forms the fp32 ``dW_acc`` during the loss forward, in the peer-visible buffer;
   as soon as its last dW contraction is done the dense exchange starts on the side stream, i.e. behind the last dX
   contraction and, in a full model, behind the whole decoder backward.  When the trainer did not announce its loss
   multiplier (``loss_fn.grad_hint``) the exchange waits for ``backward`` instead, where the multiplier is known;
3. **embedding part** — it only exists once the decoder backward has produced the gradient at the embedding output,
   so it cannot join ``dW_acc`` before (2): the embedding backward publishes this rank's rows summed per distinct id
   and every rank adds all ranks' rows to its ``.grad`` rows (sparse exchange: U x D x 2 bytes per rank).

``exchange="nccl"`` is the plain baseline: bf16 ``all_reduce`` of THIS micro-batch's contribution after the embedding
backward (into ``.grad`` directly when it did not exist, else through a scratch buffer that is then added, so earlier
micro-batches — already global sums — are never reduced twice).
"""
from __future__ import annotations

import os
import warnings

import torch
import torch.distributed as dist
from torch import nn

from . import ops


class TiedGradSync:
    def __init__(self, weight: nn.Parameter, process_group, exchange: str | None = None, max_rows: int = 65536):
        if not dist.is_initialized():
            raise RuntimeError("ssi_b200.sync: torch.distributed is not initialised")
        if weight.dtype != torch.bfloat16 or weight.dim() != 2 or not weight.is_cuda:
            raise ValueError("ssi_b200.sync: the tied weight must be a bf16 [V, D] CUDA tensor")
        self.weight = weight
        self.pg = process_group
        self.world = dist.get_world_size(process_group)
        self.rank = dist.get_rank(process_group)
        self.device = weight.device
        asked = exchange or os.environ.get("SSI_B200_EXCHANGE")
        self.mode = asked or "peer"
        if self.mode not in ("peer", "nccl", "nvls"):
            raise ValueError(f"ssi_b200: unknown exchange {self.mode!r} (expected 'peer', 'nvls' or 'nccl')")
        # "nvls" = the peer exchange with the in-switch reduction (multimem.ld_reduce) for the dense part
        nvls = True if self.mode == "nvls" else None
        if self.mode == "nvls":
            self.mode = "peer"
        self.peer = None
        if self.mode == "peer":
            from .peer import TiedGradExchange

            try:
                self.peer = TiedGradExchange(weight.size(0), weight.size(1), process_group, self.device, max_rows,
                                             nvls=nvls)
            except RuntimeError as e:
                if asked in ("peer", "nvls"):
                    raise  # asked for explicitly
                # the failure was agreed on by all ranks (TiedGradExchange), so all of them land here together
                warnings.warn(f"{e}; using the NCCL all-reduce of the tied gradient instead")
                self.mode = "nccl"
        if self.peer is not None:
            # SSI_B200_PEER_SERIAL=1 (diagnostics): run the exchanges on the main stream instead of beside it
            self.serial = os.environ.get("SSI_B200_PEER_SERIAL") == "1"
            self.side = torch.cuda.current_stream(self.device) if self.serial else torch.cuda.Stream(self.device)
            self.dw_ready, self.counts_ready, self.local_ready, self.dense_done = (torch.cuda.Event() for _ in range(4))
            for e in (self.dw_ready, self.counts_ready, self.local_ready, self.dense_done):
                e.record(torch.cuda.current_stream(self.device))  # creates the CUDA handle
        self._scratch = None          # nccl mode: this micro-batch's contribution when .grad already holds earlier ones
        self._target = None           # nccl mode: where this micro-batch's LM-head gradient went
        self._dense_done = False      # peer mode: this micro-batch's dense exchange is enqueued
        self.grad_hint_mismatch: torch.Tensor | None = None
        # SSI_B200_TIMELINE=1 (diagnostics): CUDA events at the seams of every micro-batch, summarised by close()
        self._tl = [] if (os.environ.get("SSI_B200_TIMELINE") == "1" and self.peer is not None) else None
        self._tl_cur = None
        self.fold = None              # ssi_b200.fold.TiedGradFold (install(fold_scale_grads=True)), peer mode only
        self._rows = 0                # token rows of the running micro-batch (sizes the fold scratch)

    # --------------------------------------------------------------------------------------------------- counts
    @property
    def dw_acc(self) -> torch.Tensor | None:
        return self.peer.dw_acc if self.peer is not None else None

    def global_counts(self, local_counts: torch.Tensor):
        """``(global_counts int64 [2], hint fp32 [1], event or None)``: the local {unshifted, shifted} valid-token
        counts summed over ranks and ``float(global_counts[0])``.  Peer mode fills them on the side stream; the
        returned event says when (``ops.lmhead_ce_fwd_grad(pre_grad_event=...)`` waits for it on the device)."""
        if self.peer is not None:
            main = torch.cuda.current_stream(self.device)
            if self._tl is not None:
                self._tl_cur = {"start": self._mark(main)}
            self.local_ready.record(main)
            self.side.wait_event(self.local_ready)
            with torch.cuda.stream(self.side):
                gcounts, hint = self.peer.counts(local_counts)
                self.counts_ready.record(self.side)
            if not self.serial:  # each tensor is allocated under one stream and used on the other
                local_counts.record_stream(self.side)
                gcounts.record_stream(main)
                hint.record_stream(main)
            return gcounts, hint, self.counts_ready
        gcounts = local_counts.clone()
        dist.all_reduce(gcounts, op=dist.ReduceOp.SUM, group=self.pg)
        return gcounts, gcounts[0:1].to(torch.float32), None

    # --------------------------------------------------------------------------------------------- LM-head part
    def _grad_buffer(self) -> bool:
        """Make sure ``weight.grad`` exists; returns True when it already held earlier micro-batches."""
        had = self.weight.grad is not None
        if not had:
            self.weight.grad = torch.empty_like(self.weight)
        return had

    def _dense(self, grad_out: torch.Tensor, hint: torch.Tensor, after_event: torch.cuda.Event | None) -> None:
        main = torch.cuda.current_stream(self.device)
        accumulate = self._grad_buffer()
        if self.fold is not None:
            self.fold.ensure(self._rows)  # allocate on the main stream, use on the side stream
        if after_event is not None:
            self.side.wait_event(after_event)
        else:
            self.side.wait_stream(main)
        with torch.cuda.stream(self.side):
            # the scale_grads / grad-norm fold rides in the reduce (the side stream is ordered after `hint`)
            ps, fws = self.fold.dense_args(hint, self._rows) if self.fold is not None else (None, None)
            if self._tl_cur is not None:
                self._tl_cur["dw_done"] = self._mark(self.side)
            self.peer.exchange(self.weight.grad, grad_out, hint, accumulate=accumulate, post_scale=ps, fold_ws=fws)
            if self._tl_cur is not None:
                self._tl_cur["dense_done"] = self._mark(self.side)
            self.dense_done.record(self.side)
        self._dense_done = True

    def before_dw_acc_write(self) -> None:
        """Call on the main stream before anything overwrites the peer-visible ``dw_acc``: the previous dense exchange
        (side stream) reads it here AND on every peer; its gather step ends only after all peers reduced their slice,
        i.e. finished reading, so waiting for the event behind it is enough for both.  A micro-batch with ONE loss
        backward never waits (the embedding backward already joined the streams); two ``compute_cross_entropy``
        backwards per micro-batch do."""
        if self.peer is not None and not self.serial:
            torch.cuda.current_stream(self.device).wait_event(self.dense_done)

    def after_forward(self, hint: torch.Tensor, announced: bool, n_rows: int = 0) -> None:
        """Called right after the single-pass forward was enqueued (it recorded ``dw_ready`` behind its last dW
        contraction).  With an announced multiplier the dense exchange starts now, behind the last dX contraction."""
        self._dense_done = False
        self._rows = n_rows
        if self.peer is not None and announced:
            self._dense(hint, hint, self.dw_ready)

    def in_backward(self, dX: torch.Tensor, dW_acc: torch.Tensor, g: torch.Tensor, hint: torch.Tensor,
                    formed_outside: bool = False) -> None:
        """Backward of the loss: dX is rescaled in place; the LM-head gradient goes to the exchange.
        ``formed_outside``: this loss's forward did not go through :meth:`after_forward` (a ``compute_cross_entropy``
        chunk): its gradient was just copied into the exchange buffer and needs an exchange of its own, however many
        such backwards a micro-batch has."""
        if self.peer is not None:
            self._rows = max(self._rows, dX.size(0))
            ops.lmhead_ce_apply_grad(dX, dW_acc, g, hint, skip_dW=True)
            if self._dense_done and not formed_outside:
                # the exchange already ran with the announced multiplier: leave a wrong announcement visible
                # (OR-ed over the micro-batches of an accumulation window until check() reads it)
                m = (g != hint)
                self.grad_hint_mismatch = m if self.grad_hint_mismatch is None else (self.grad_hint_mismatch | m)
            else:
                self._dense(g, hint, None)
            return
        # nccl: this micro-batch's contribution goes where it can be all-reduced alone
        if self.weight.grad is None:
            self.weight.grad = torch.empty_like(self.weight)
            self._target = self.weight.grad
        else:
            if self._scratch is None:
                self._scratch = torch.empty_like(self.weight)
            self._target = self._scratch
        ops.lmhead_ce_apply_grad(dX, dW_acc, g, hint, dW=self._target, accumulate_dW=False)

    def recompute_target(self) -> tuple[torch.Tensor, bool]:
        """Recompute mode (its backward writes bf16 dW directly): where to write and whether to accumulate."""
        if self.weight.grad is None:
            self.weight.grad = torch.empty_like(self.weight)
            self._target = self.weight.grad
        else:
            if self._scratch is None:
                self._scratch = torch.empty_like(self.weight)
            self._target = self._scratch
        return self._target, False

    # -------------------------------------------------------------------------------------------- embedding part
    def embedding_backward(self, dH: torch.Tensor, sort_ws: torch.Tensor) -> None:
        """The last contribution of a micro-batch to the tied gradient (backward of ``tok_embeddings``)."""
        main = torch.cuda.current_stream(self.device)
        if self.peer is not None:
            if self._tl_cur is not None:
                self._tl_cur["bwd_done"] = self._mark(main)
            if not self.serial:
                main.wait_stream(self.side)  # the dense exchange wrote .grad; peers are done with last step's rows
            if self.weight.grad is None:
                self.weight.grad = torch.zeros_like(self.weight)
            ps, fws = self.fold.sparse_args(dH.size(0)) if self.fold is not None else (None, None)
            self.peer.embedding_exchange(dH, sort_ws, self.weight.grad, post_scale=ps, fold_ws=fws)
            self._dense_done = False
            if self._tl_cur is not None:
                self._tl_cur["end"] = self._mark(main)
                if len(self._tl_cur) == 5:
                    self._tl.append(self._tl_cur)
                self._tl_cur = None
            return
        target = self._target
        if target is None:  # no LM-head backward in this micro-batch
            if self.weight.grad is None:
                self.weight.grad = torch.zeros_like(self.weight)
                target = self.weight.grad
            else:
                if self._scratch is None:
                    self._scratch = torch.empty_like(self.weight)
                target = self._scratch
                target.zero_()
        ops.embedding_grad_scatter_sorted(dH, sort_ws, target)
        dist.all_reduce(target, op=dist.ReduceOp.SUM, group=self.pg)
        if target is not self.weight.grad:
            ops.grad_add(self.weight.grad, target)
        self._target = None

    def check(self, confirm_multiplier: bool = False) -> None:
        """Raise if an exchange timed out (a pinned host word: no device synchronisation).  ``confirm_multiplier``:
        also raise if the dense exchange ran with a multiplier that backward did not confirm (reads a device flag:
        synchronises; meant for the end of a run or an accumulation boundary)."""
        if self.peer is not None:
            self.peer.check()
        m = self.grad_hint_mismatch if confirm_multiplier else None
        if m is not None:
            self.grad_hint_mismatch = None
        if m is not None and bool(m.any().item()):
            raise RuntimeError("ssi_b200.sync: the dense exchange ran with the announced loss multiplier, but "
                               "backward received a different upstream gradient; announce the right one "
                               "(loss_fn.grad_hint) or none at all")

    @staticmethod
    def _mark(stream) -> torch.cuda.Event:
        e = torch.cuda.Event(enable_timing=True)
        e.record(stream)
        return e

    def timeline(self, last: int = 10) -> dict | None:
        """Mean milliseconds, over the last micro-batches, from the start of the loss forward to: the last dW
        contraction done (side stream released), the dense exchange done, dX returned (loss backward done), the end of
        the embedding exchange.  Diagnostics (``SSI_B200_TIMELINE=1``); synchronises."""
        if not self._tl:
            return None
        torch.cuda.synchronize(self.device)
        rows = self._tl[-last:]
        keys = ("dw_done", "dense_done", "bwd_done", "end")
        return {k: sum(r["start"].elapsed_time(r[k]) for r in rows) / len(rows) for k in keys}

    def close(self) -> None:
        """Release the peer exchange buffers (collective: every rank must call it)."""
        if self._tl:
            import sys

            print(f"ssi_b200 timeline rank {self.rank}: {self.timeline()}", file=sys.stderr, flush=True)
        if self.peer is not None:
            self.peer.close()
            self.peer = None
