"""Data-parallel plumbing for the path: the token batch is sharded by rows across ranks (what the reference's
``DistributedSampler`` scaffold would do, ssi/data/__init__.py:82-83), W is replicated, and the only exchange
is (1) a 16-byte all-reduce of the token counts before backward, so that every rank scales its gradient by the
GLOBAL number of valid tokens, and (2) an all-reduce (sum) of the tied LM-head/embedding gradient ``[V, D]``.
Semantics follow torchtune's distributed recipe (tt/recipes/full_finetune_distributed.py:844-851,867): losses and
token counts are summed over ranks, gradients are those of the global mean.

One process per GPU; ``torch.distributed`` with the NCCL backend over NVLink 5 / NVSwitch on B200, gloo in the
CPU tests.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def global_counts(counts: torch.Tensor, group=None) -> torch.Tensor:
    """Sum the int64 ``[2]`` (unshifted, shifted) valid-token counts over ranks; returns a new tensor."""
    out = counts.clone()
    dist.all_reduce(out, op=dist.ReduceOp.SUM, group=group)
    return out


def shard_rows(n_rows: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous row range [begin, end) of ``rank`` when ``n_rows`` sequences are split over ``world`` ranks
    (earlier ranks take the remainder, like torch.chunk / DistributedSampler without padding)."""
    base, rem = divmod(n_rows, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def allreduce_tied_grad(grad: torch.Tensor, group=None, bucket_rows: int = 0) -> None:
    """In-place sum of the tied weight gradient over ranks.  ``bucket_rows`` > 0 splits the ``[V, D]`` buffer into
    row buckets issued back to back (NVSwitch makes cost per byte uniform, so buckets only matter for overlap)."""
    if bucket_rows <= 0 or bucket_rows >= grad.size(0):
        dist.all_reduce(grad, op=dist.ReduceOp.SUM, group=group)
        return
    for r0 in range(0, grad.size(0), bucket_rows):
        dist.all_reduce(grad[r0:r0 + bucket_rows], op=dist.ReduceOp.SUM, group=group)


def allreduce_loss(loss_batch: torch.Tensor, group=None) -> torch.Tensor:
    """Sum of the per-rank ``loss * num_tokens`` contributions (tt/recipes/full_finetune_distributed.py:847)."""
    out = loss_batch.detach().clone().reshape(1)
    dist.all_reduce(out, op=dist.ReduceOp.SUM, group=group)
    return out[0]
