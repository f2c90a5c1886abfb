"""GPU tests of the data-parallel exchange of the tied gradient over peer memory (csrc/dp_exchange.cuh,
ssi_b200.peer, PathStep(exchange="peer")): the fp32 embedding scatter-add, the reduce + gather kernels and the
side-stream overlap, against the fp32 oracle on the CONCATENATED batch (tolerances as in test_gpu_parity.py).

world = 1 runs in-process on any GPU box (the kernels' own-rank path); world = 2 spawns one process per GPU over
NCCL and needs two GPUs (skipped otherwise; `gpurun --gpus 2`)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import reference_path as oracle
from test_gpu_parity import LOSS_RTOL, assert_grad_close

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _inputs(B=4, S=160, V=5002, D=256, cfg="c2", seed=5):
    from ssi_b200 import synth

    return synth.make_inputs(synth.CONFIGS[cfg], batch=B, seq=S, vocab=V, hidden=D, seed=synth.SEED + seed)


def test_scatter_f32_matches_oracle():
    from ssi_b200 import ops

    inp = _inputs()
    V, D = inp["W"].shape
    base = torch.randn(V, D)
    acc = base.clone().cuda()
    ops.embedding_grad_scatter_f32(inp["dH0"].reshape(-1, D).cuda(), inp["tokens"].reshape(-1).cuda(), acc)
    torch.cuda.synchronize()
    ref = base + oracle.embedding_grad(inp["tokens"], inp["dH0"], V)
    assert torch.allclose(acc.cpu(), ref, rtol=1e-5, atol=1e-6)


def test_peer_exchange_world1_matches_oracle_and_plain_path():
    from ssi_b200.step import PathStep

    port = _free_port()
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=0, world_size=1)
    try:
        inp = _inputs()
        dev = torch.device("cuda", 0)
        args = [inp[k].to(dev) for k in ("tokens", "labels", "X", "dH0")]
        step = PathStep(inp["W"].to(dev), process_group=dist.group.WORLD, exchange="peer")
        plain = PathStep(inp["W"].to(dev))
        outs = []
        for _ in range(3):  # several epochs through the same buffers; results must not drift
            step.zero_grad()
            loss_batch, dX, counts = step(*args)
            torch.cuda.synchronize()
            outs.append((loss_batch.clone(), dX.clone(), step.weight.grad.clone()))
        assert not bool(step.loss_fn.grad_hint_mismatch.any())
        ref_loss, n_iter, ref_dH, ref_dW = oracle.train_step_path(inp["tokens"], inp["labels"], inp["X"], inp["dH0"],
                                                                  inp["W"])
        want = ref_loss.item() * n_iter
        assert abs(outs[0][0].item() - want) <= LOSS_RTOL * abs(want)
        assert_grad_close(outs[0][1].reshape(ref_dH.shape), ref_dH, "dH")
        assert_grad_close(outs[0][2], ref_dW, "dW(head+embedding), peer exchange")
        for o in outs[1:]:
            assert torch.equal(o[1], outs[0][1]) and torch.equal(o[2], outs[0][2])
        # gradient accumulation across micro-batches: second call adds to the existing .grad
        step(*args)
        torch.cuda.synchronize()
        assert_grad_close(step.weight.grad, 2 * ref_dW, "dW accumulated over two micro-batches")
        plain(*args)
        torch.cuda.synchronize()
        assert_grad_close(plain.weight.grad, outs[0][2].float(), "peer vs plain path")
        step.close()
    finally:
        dist.destroy_process_group()


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from ssi_b200 import dist as ddist
        from ssi_b200.step import PathStep

        inp = _inputs(B=6, S=200, V=9001, D=512, cfg="c3", seed=9)
        b0, b1 = ddist.shard_rows(inp["tokens"].size(0), rank, world)
        args = [inp[k][b0:b1].contiguous().to(dev) for k in ("tokens", "labels", "X", "dH0")]
        res = {}
        for ex in ("peer", "nccl"):
            step = PathStep(inp["W"].to(dev), process_group=dist.group.WORLD, exchange=ex)
            for _ in range(2):
                step.zero_grad()
                loss_batch, dX, counts = step(*args)
            torch.cuda.synchronize()
            total = ddist.allreduce_loss(loss_batch, dist.group.WORLD)
            res[ex] = (total.item(), dX.cpu(), step.weight.grad.cpu())
            step.close()
        # every rank must hold the bit-identical summed gradient
        g = res["peer"][2].cuda().view(torch.int16).to(torch.int32)
        lo, hi = g.clone(), g.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        same = bool((lo == hi).all().item())
        if rank == 0:
            ref_loss, n_iter, ref_dH, ref_dW = oracle.train_step_path(inp["tokens"], inp["labels"], inp["X"],
                                                                      inp["dH0"], inp["W"])
            want = ref_loss.item() * n_iter
            out["same_on_all_ranks"] = same
            out["loss_ok"] = abs(res["peer"][0] - want) <= LOSS_RTOL * abs(want)
            try:
                assert_grad_close(res["peer"][2], ref_dW, "dW peer")
                assert_grad_close(res["nccl"][2], ref_dW, "dW nccl")
                assert_grad_close(res["peer"][1].reshape(ref_dH[b0:b1].shape), ref_dH[b0:b1], "dH")
                out["grads_ok"] = True
            except AssertionError as e:  # pragma: no cover
                out["grads_ok"] = False
                out["why"] = str(e)
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_peer_exchange_world2_matches_single_process_oracle():
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    assert out.get("same_on_all_ranks") is True
    assert out.get("loss_ok") is True
    assert out.get("grads_ok") is True, out.get("why")
