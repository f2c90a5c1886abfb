"""GPU tests of the data-parallel exchange of the tied gradient (csrc/dp_exchange.cuh, ssi_b200.peer,
ssi_b200.sync) in the CAUSAL order of a training step -- token counts through the peers' ctrl blocks, the LM-head part
behind the last dX contraction, the embedding part (compact rows) after the embedding backward -- through PathStep and
through install() on a tiny model whose embedding gradient arrives by autograd, against the fp32 oracle on the
CONCATENATED batch (tolerances as in test_gpu_parity.py).

world = 1 runs in-process on any GPU box (the kernels' own-rank path); world = 2 spawns one process per GPU over
NCCL and needs two GPUs (skipped otherwise; `gpurun --gpus 2`)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import reference_path as oracle
from test_gpu_parity import DW_ELEM_ROUNDINGS, LOSS_RTOL, assert_grad_close

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
        assert_grad_close(outs[0][2], ref_dW, "dW(head+embedding), peer exchange", elem_roundings=DW_ELEM_ROUNDINGS)
        for o in outs[1:]:
            assert torch.equal(o[1], outs[0][1]) and torch.equal(o[2], outs[0][2])
        # gradient accumulation across micro-batches: second call adds to the existing .grad
        step(*args)
        torch.cuda.synchronize()
        assert_grad_close(step.weight.grad, 2 * ref_dW, "dW accumulated over two micro-batches",
                          elem_roundings=DW_ELEM_ROUNDINGS)
        plain(*args)
        torch.cuda.synchronize()
        assert_grad_close(plain.weight.grad, outs[0][2].float(), "peer vs plain path")
        # two-phase form of the sparse exchange (owner-reduce + apply; default only for world > 4), forced here
        from ssi_b200 import _lib

        _lib.set_option("sparse_two_phase", 1)
        try:
            two = []
            for _ in range(2):
                step.zero_grad()
                step(*args)
                torch.cuda.synchronize()
                two.append(step.weight.grad.clone())
        finally:
            _lib.set_option("sparse_two_phase", -1)
        assert torch.equal(two[0], two[1])
        assert_grad_close(two[0], ref_dW, "dW, two-phase sparse exchange", elem_roundings=DW_ELEM_ROUNDINGS)
        step.close()
    finally:
        dist.destroy_process_group()


class TinyDecoder(torch.nn.Module):
    """The reference model's shape of interface (tok_embeddings / norm-free / output, chunked output list,
    tt/torchtune/modules/transformer.py:499-502,628) with one mixing parameter, so that the embedding gradient reaches
    ``tok_embeddings`` through autograd AFTER the loss backward -- the causal order of a real step."""

    def __init__(self, V, D):
        super().__init__()
        self.tok_embeddings = torch.nn.Embedding(V, D, dtype=torch.bfloat16)
        self.output = None
        self.num_output_chunks = 0
        self.mix = torch.nn.Parameter(torch.tensor(0.5, dtype=torch.bfloat16))

    def set_num_output_chunks(self, n):
        self.num_output_chunks = n

    def forward(self, tokens, extra):
        h = self.tok_embeddings(tokens) * self.mix + extra
        return [self.output(c) for c in h.chunk(self.num_output_chunks, dim=1)]


def _oracle_tiny(inp, mbs):
    """fp32 CPU oracle of TinyDecoder on the CONCATENATED batch, micro-batches accumulated."""
    Wf = inp["W"].float().clone().requires_grad_(True)
    mix = torch.tensor(0.5, requires_grad=True)
    total = 0.0
    for sl in mbs:
        tok, lab, X = inp["tokens"][sl], inp["labels"][sl], inp["X"][sl]
        n_iter = oracle.count_valid(lab)
        h = torch.nn.functional.embedding(tok, Wf) * mix + X.float()
        loss = oracle.lmhead_ce(h, Wf, oracle.shift_labels(lab)) * n_iter
        loss.backward()
        total += loss.item()
    return total, Wf.grad, mix.grad


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        import ssi_b200
        from ssi_b200 import dist as ddist
        from ssi_b200 import ops
        from ssi_b200.step import PathStep

        pg = dist.group.WORLD
        # ---------------- (1) PathStep: the bench's flow, both exchanges, against the oracle on the whole batch
        inp = _inputs(B=6, S=200, V=9001, D=512, cfg="c3", seed=9)
        b0, b1 = ddist.shard_rows(inp["tokens"].size(0), rank, world)
        args = [inp[k][b0:b1].contiguous().to(dev) for k in ("tokens", "labels", "X", "dH0")]
        res = {}
        for ex in ("peer", "nccl"):
            step = PathStep(inp["W"].to(dev), process_group=pg, exchange=ex)
            for _ in range(2):
                step.zero_grad()
                loss_batch, dX, counts = step(*args)
            torch.cuda.synchronize()
            total = ddist.allreduce_loss(loss_batch, pg)
            res[ex] = (total.item(), dX.cpu(), step.weight.grad.cpu())
            # a second micro-batch WITHOUT zero_grad: only the new contribution may be summed over ranks
            step(*args)
            torch.cuda.synchronize()
            res[ex + "_acc"] = step.weight.grad.cpu()
            # eval after a training step: this rank's own mean, not a stale global count
            ev = step.forward_only(args[2], args[1])
            res[ex + "_eval"] = ev.item()
            step.close()
        # scale_grads + grad norm folded into the peer reduce / gather and the sparse embedding exchange (8f-3)
        step = PathStep(inp["W"].to(dev), process_group=pg, exchange="peer", fold_scale_grads=True)
        for _ in range(2):
            step.zero_grad()
            step(*args)
            fold_sq = step.fold.sqnorm()
        torch.cuda.synchronize()
        res["fold"] = (fold_sq.item(), None, step.weight.grad.cpu())
        step.close()
        # two-phase sparse exchange (owner-reduce + apply), forced at world 2, with the fold riding along
        from ssi_b200 import _lib as _l

        _l.set_option("sparse_two_phase", 1)
        try:
            step = PathStep(inp["W"].to(dev), process_group=pg, exchange="peer", fold_scale_grads=True)
            for _ in range(2):
                step.zero_grad()
                step(*args)
                tp_sq = step.fold.sqnorm()
            torch.cuda.synchronize()
            res["twophase"] = (tp_sq.item(), None, step.weight.grad.cpu())
            step.close()
        finally:
            _l.set_option("sparse_two_phase", -1)
        # the in-switch reduction (NVLS, multimem.ld_reduce) for the dense part, where the box has multicast support
        try:
            step = PathStep(inp["W"].to(dev), process_group=pg, exchange="nvls")
        except RuntimeError as e:
            step, res["nvls_why"] = None, str(e)
        if step is not None:
            grads = []
            for _ in range(2):
                step.zero_grad()
                step(*args)
                torch.cuda.synchronize()
                grads.append(step.weight.grad.clone())
            res["nvls"] = (None, None, grads[1].cpu())
            res["nvls_repeatable"] = bool(torch.equal(grads[0], grads[1]))
            step.close()
        # every rank must hold the bit-identical summed gradient
        same = {}
        for ex in ("peer", "nccl", "fold", "twophase") + (("nvls",) if "nvls" in res else ()):
            g = res[ex][2].cuda().view(torch.int16).to(torch.int32)
            lo, hi = g.clone(), g.clone()
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX)
            same[ex] = bool((lo == hi).all().item())

        # ---------------- (2) through install() on a tiny model, two micro-batches, multiplier NOT announced
        V, D, B, S = 3001, 256, 4, 96
        tin = _inputs(B=2 * B, S=S, V=V, D=D, cfg="c2", seed=13)
        model = TinyDecoder(V, D).to(dev)
        with torch.no_grad():
            model.tok_embeddings.weight.copy_(tin["W"])
        tiny = {}
        for ex in ("peer", "nccl"):
            model.zero_grad(set_to_none=True)
            loss_fn = ssi_b200.install(model, process_group=pg, exchange=ex, max_tokens_per_rank=B * S)
            tot = 0.0
            for mb in range(2):  # micro-batch mb = global rows [mb*B, (mb+1)*B), split over the ranks
                r0, r1 = ddist.shard_rows(B, rank, world)
                sl = slice(mb * B + r0, mb * B + r1)
                tok, lab, X = (tin[k][sl].to(dev) for k in ("tokens", "labels", "X"))
                shifted, cnt = ops.label_shift_count(lab.contiguous())
                loss = loss_fn(model(tok, X), shifted)  # local loss sum / GLOBAL valid count
                n_iter = ddist.global_counts(cnt, pg)[0]  # the trainer's own all-reduced multiplier
                lb = loss * n_iter
                lb.backward()
                tot += ddist.allreduce_loss(lb, pg).item()
            torch.cuda.synchronize()
            mixg = model.mix.grad.float().clone()
            dist.all_reduce(mixg)  # an ordinary parameter: the trainer's (DDP's) job, summed here by hand
            tiny[ex] = (tot, model.tok_embeddings.weight.grad.cpu().clone(), mixg.item())
            loss_fn.sync.check(confirm_multiplier=True)
            loss_fn.sync.close()

        if rank == 0:
            ref_loss, n_iter, ref_dH, ref_dW = oracle.train_step_path(inp["tokens"], inp["labels"], inp["X"],
                                                                      inp["dH0"], inp["W"])
            want = ref_loss.item() * n_iter
            out["same_on_all_ranks"] = same["peer"] and same["nccl"] and same["fold"] and same["twophase"]
            out["loss_ok"] = (abs(res["peer"][0] - want) <= LOSS_RTOL * abs(want)
                              and abs(res["nccl"][0] - want) <= LOSS_RTOL * abs(want))
            lab_loc = inp["labels"][b0:b1]
            ev_want = oracle.lmhead_ce(inp["X"][b0:b1].float(), inp["W"].float(), oracle.shift_labels(lab_loc)).item()
            out["eval_ok"] = all(abs(res[k + "_eval"] - ev_want) <= LOSS_RTOL * abs(ev_want) for k in ("peer", "nccl"))
            try:
                assert_grad_close(res["peer"][2], ref_dW, "dW peer", elem_roundings=DW_ELEM_ROUNDINGS)
                assert_grad_close(res["nccl"][2], ref_dW, "dW nccl", elem_roundings=DW_ELEM_ROUNDINGS)
                assert_grad_close(res["peer_acc"], 2 * ref_dW, "dW peer, two micro-batches accumulated",
                                  elem_roundings=DW_ELEM_ROUNDINGS)
                assert_grad_close(res["nccl_acc"], 2 * ref_dW, "dW nccl, two micro-batches accumulated",
                                  elem_roundings=DW_ELEM_ROUNDINGS)
                assert_grad_close(res["peer"][1].reshape(ref_dH[b0:b1].shape), ref_dH[b0:b1], "dH")
                assert_grad_close(res["fold"][2], ref_dW / n_iter, "dW peer, scale_grads folded",
                                  elem_roundings=DW_ELEM_ROUNDINGS)
                assert_grad_close(res["twophase"][2], ref_dW / n_iter, "dW peer, two-phase sparse exchange, folded",
                                  elem_roundings=DW_ELEM_ROUNDINGS)
                tp_own = res["twophase"][2].double().pow(2).sum().item()
                assert abs(res["twophase"][0] - tp_own) <= 5e-5 * tp_own, ("two-phase sqnorm", res["twophase"][0])
                if "nvls" in res:
                    assert same["nvls"] and res["nvls_repeatable"], "NVLS reduce: not identical on all ranks / runs"
                    assert_grad_close(res["nvls"][2], ref_dW, "dW peer + NVLS reduce", elem_roundings=DW_ELEM_ROUNDINGS)
                    out["nvls"] = "tested"
                else:
                    out["nvls"] = "unavailable: " + res.get("nvls_why", "?")[:200]
                own_sq = res["fold"][2].double().pow(2).sum().item()
                assert abs(res["fold"][0] - own_sq) <= 5e-5 * own_sq, ("folded sqnorm", res["fold"][0], own_sq)
                t_loss, t_dW, t_mix = _oracle_tiny(tin, [slice(0, B), slice(B, 2 * B)])
                for ex in ("peer", "nccl"):
                    assert abs(tiny[ex][0] - t_loss) <= 2e-3 * abs(t_loss), (ex, tiny[ex][0], t_loss)
                    # nccl baseline: two micro-batches x two ranks summed in bf16 (one rounding per hop and add)
                    assert_grad_close(tiny[ex][1], t_dW, f"tied dW through install(), {ex}",
                                      roundings=2.0 if ex == "nccl" else 1.0, elem_roundings=DW_ELEM_ROUNDINGS)
                    assert abs(tiny[ex][2] - t_mix.item()) <= 2e-2 * abs(t_mix.item()) + 1e-3, (ex, tiny[ex][2])
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
    assert out.get("eval_ok") is True
    assert out.get("grads_ok") is True, out.get("why")
    print("NVLS dense reduce:", out.get("nvls"))


def _timeout_worker(rank, world, port, out):
    """Rank 1 arrives 3 s late at an exchange whose wait limit is 1 s."""
    import time

    from ssi_b200 import _lib
    from ssi_b200.peer import TiedGradExchange

    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                            device_id=dev)
    try:
        V, D = 1024, 128
        _lib.set_option("dp_timeout_s", 1)
        px = TiedGradExchange(V, D, max_rows=256, group=dist.group.WORLD, device=dev, nvls=False)
        one = torch.ones(1, dtype=torch.float32, device=dev)

        def step():
            px.dw_acc.fill_(float(rank + 1))
            dW = torch.zeros(V, D, dtype=torch.bfloat16, device=dev)
            px.exchange(dW, one, one)
            torch.cuda.synchronize()
            return dW

        ok_first = bool((step() == 3).all().item())  # 1 + 2: an ordinary exchange
        dist.barrier()
        if rank == 1:
            time.sleep(3.0)
        late = step()
        # the rank that waited gave up and marked BOTH ranks; the late rank must not come away with finite numbers
        nan_everywhere = bool(torch.isnan(late.float()).all().item())
        flagged = px.timed_out()
        raised = False
        try:
            px.check()
        except RuntimeError:
            raised = True
        _lib.set_option("dp_timeout_s", 600)
        px.reset()
        ok_after = bool((step() == 3).all().item()) and not px.timed_out()
        out[rank] = (ok_first, nan_everywhere, flagged, raised, ok_after)
        px.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_a_wait_that_times_out_poisons_every_rank_and_reset_recovers():
    """the device-side wait used to give up after a hard-coded 20 s, marked only the rank that
    waited (the late rank then read buffers being rewritten and kept finite, wrong gradients), and never recovered.
    Now the limit is an option (default 10 min), a time-out marks every rank, and reset() is the collective way back."""
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_timeout_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    for rank in (0, 1):
        assert out.get(rank) == (True, True, True, True, True), (rank, out.get(rank))


def _two_chunk_worker(rank, world, port, out):
    """Two ``compute_cross_entropy`` chunks per micro-batch and rank (what ``compile_loss`` drives), peer exchange."""
    import ssi_b200
    from ssi_b200 import ops

    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    if world == 1:  # in-process on a one-GPU box: the exchange kernels' own-rank path
        dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=0, world_size=1)
    else:
        dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                                device_id=dev)
    try:
        V, D, B, S = 3001, 256, 4, 96
        tin = _inputs(B=B, S=S, V=V, D=D, cfg="c2", seed=17)
        n_iter = oracle.count_valid(tin["labels"])
        n_shift = oracle.count_valid(oracle.shift_labels(tin["labels"]))
        model = TinyDecoder(V, D).to(dev)
        with torch.no_grad():
            model.tok_embeddings.weight.copy_(tin["W"])
        loss_fn = ssi_b200.install(model, process_group=dist.group.WORLD, exchange="peer", max_tokens_per_rank=B * S)
        model.set_num_output_chunks(2)
        b = B // world
        sl = slice(rank * b, (rank + 1) * b)
        grads = []
        for _ in range(2):  # twice: the second pass runs on top of the first one's buffers and must repeat it
            model.zero_grad(set_to_none=True)
            tok, lab, X = (tin[k][sl].to(dev) for k in ("tokens", "labels", "X"))
            shifted, _ = ops.label_shift_count(lab.contiguous())
            chunks = model(tok, X)
            total = sum(loss_fn.compute_cross_entropy(hc, lc.contiguous())
                        for hc, lc in zip(chunks, shifted.chunk(2, dim=1)))
            (total * (n_iter / n_shift)).backward()
            torch.cuda.synchronize()
            grads.append(model.tok_embeddings.weight.grad.clone())
        loss_fn.sync.check()
        loss_fn.sync.close()
        tot = (total.detach() * (n_iter / n_shift)).reshape(1).clone()
        if world > 1:
            dist.all_reduce(tot)
        if rank == 0:
            t_loss, t_dW, _ = _oracle_tiny(tin, [slice(0, B)])
            ok, why = True, ""
            try:
                assert abs(tot.item() - t_loss) <= 2e-3 * abs(t_loss), (tot.item(), t_loss)
                assert torch.equal(grads[0], grads[1]), "not repeatable"
                assert_grad_close(grads[1], t_dW, "tied dW, two compute_cross_entropy chunks per rank",
                                  elem_roundings=DW_ELEM_ROUNDINGS)
            except AssertionError as e:  # pragma: no cover
                ok, why = False, str(e)
            out["ok"], out["why"] = ok, why
    finally:
        dist.destroy_process_group()


def test_two_compute_cross_entropy_backwards_per_micro_batch_world1():
    """The same on one GPU (world 1): both chunks' gradients must reach the tied weight's .grad."""
    out = {}
    _two_chunk_worker(0, 1, _free_port(), out)
    assert out.get("ok") is True, out.get("why")


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_compute_cross_entropy_backwards_per_micro_batch_are_both_exchanged():
    """every chunk's LM-head gradient is copied into the one peer-visible fp32 buffer and exchanged
    from there; the copy of chunk k+1 must wait for the exchange of chunk k (local and remote readers), and every
    chunk needs its own exchange (the second one used to be taken for "already exchanged in forward")."""
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_two_chunk_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    assert out.get("ok") is True, out.get("why")
