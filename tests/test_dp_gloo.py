"""world_size-2 gloo test of the data-parallel host logic (no GPU): sharding the batch by sequences, summing the
token counts, scaling every rank's gradient by the GLOBAL counts and all-reducing the tied gradient reproduces
the single-process oracle on the concatenated batch.  The per-rank arithmetic here is the CPU oracle standing
in for the CUDA kernels (the kernels themselves are covered by the -m gpu tests)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import reference_path as oracle
from ssi_b200 import dist as ddist
from ssi_b200 import synth


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cfg = synth.CONFIGS["c2"]
        B, S, V, D = 4, 48, 300, 32
        inp = synth.make_inputs(cfg, batch=B, seq=S, vocab=V, hidden=D, seed=11)
        b0, b1 = ddist.shard_rows(B, rank, world)
        tok, lab, X, dH0 = (inp[k][b0:b1] for k in ("tokens", "labels", "X", "dH0"))
        shifted = oracle.shift_labels(lab)
        counts = torch.tensor([oracle.count_valid(lab), oracle.count_valid(shifted)], dtype=torch.int64)
        g = ddist.global_counts(counts)
        # local mean-over-local-valid loss, rescaled to the global divisor, seeded with the global token count
        hh = X.float().clone().requires_grad_(True)
        ww = inp["W"].float().clone().requires_grad_(True)
        loss_local = oracle.lmhead_ce(hh, ww, shifted) * counts[1] / g[1]
        (loss_local * g[0]).backward()
        grad = ww.grad + oracle.embedding_grad(tok, dH0, V)
        ddist.allreduce_tied_grad(grad, bucket_rows=128 if rank >= 0 else 0)
        total = ddist.allreduce_loss((loss_local * g[0]).detach())
        if rank == 0:
            ref_loss, n_iter, ref_dH, ref_dW = oracle.train_step_path(inp["tokens"], inp["labels"], inp["X"],
                                                                      inp["dH0"], inp["W"])
            out["loss_ok"] = bool(torch.allclose(total, ref_loss * n_iter, rtol=1e-5))
            out["dw_ok"] = bool(torch.allclose(grad, ref_dW, rtol=1e-4, atol=1e-6))
            out["dh_ok"] = bool(torch.allclose(hh.grad, ref_dH[b0:b1], rtol=1e-4, atol=1e-7))
            out["counts"] = g.tolist()
            out["n_iter"] = n_iter
    finally:
        dist.destroy_process_group()


def test_dp2_matches_single_process_oracle():
    mgr = mp.Manager()
    out = mgr.dict()
    port = _free_port()
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    assert out["loss_ok"] and out["dw_ok"] and out["dh_ok"], dict(out)
    assert out["counts"][0] == out["n_iter"]


def _peer_setup_worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from ssi_b200.peer import TiedGradExchange

        try:
            TiedGradExchange(300, 64, dist.group.WORLD, torch.device("cuda", 0))
            out[rank] = "no error"
        except RuntimeError as e:
            out[rank] = str(e)
        dist.barrier()  # both ranks got here: nobody hangs in a collective the other one skipped
    finally:
        dist.destroy_process_group()


def test_peer_setup_failure_is_agreed_on_by_all_ranks():
    """On a box without a GPU the exchange buffers cannot be allocated: every rank must raise the same RuntimeError
    (so that callers can fall back to NCCL together) instead of one rank raising while the other waits."""
    if torch.cuda.is_available():
        import pytest

        pytest.skip("needs a box without CUDA to provoke the set-up failure")
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_peer_setup_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    assert len(out) == 2
    for r in (0, 1):
        assert "peer-memory set-up failed" in out[r], out[r]
