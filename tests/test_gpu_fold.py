"""scale_grads + the tied weight's share of clip_grad_norm_ FOLDED into the kernels that write the tied gradient
(ssi/trainer.py:404-408, tt/torchtune/training/_grad_scaler.py:25-31), against

  * the fp32 oracle: (LM-head + embedding gradient) * scaler, and its sum of squares;
  * the two-pass form the fold replaces: the unfolded step followed by the reference's ``p.grad *= scaler`` (one more
    bf16 rounding there, so the folded result differs from it by at most one bf16 ulp almost everywhere -- and is the
    closer one to fp32).
"""
import socket

import pytest
import torch
import torch.distributed as dist

from oracle import reference_path as oracle
from test_gpu_parity import DW_ELEM_ROUNDINGS, assert_grad_close

pytestmark = pytest.mark.gpu


def _inputs(B=4, S=160, V=5002, D=256, cfg="c2", seed=21):
    from ssi_b200 import synth

    return synth.make_inputs(synth.CONFIGS[cfg], batch=B, seq=S, vocab=V, hidden=D, seed=synth.SEED + seed)


def _check_against_oracle(inp, grad, sq, micro_batches=1):
    ref_loss, n_iter, _, ref_dW = oracle.train_step_path(inp["tokens"], inp["labels"], inp["X"], inp["dH0"], inp["W"])
    scaler = 1.0 / (micro_batches * n_iter)
    want = micro_batches * ref_dW * scaler
    assert_grad_close(grad, want, "folded dW (head + embedding) * scaler", elem_roundings=DW_ELEM_ROUNDINGS)
    # the folded sum of squares is that of the bf16 values actually written
    own = grad.double().pow(2).sum().item()
    assert abs(sq - own) <= 5e-5 * own, (sq, own)
    ref_sq = want.double().pow(2).sum().item()
    assert abs(sq - ref_sq) <= 5e-3 * ref_sq, (sq, ref_sq)
    return scaler


@pytest.mark.parametrize("shape", [dict(), dict(B=3, S=333, V=9001, D=512, cfg="c3"), dict(B=2, S=64, V=777, D=64)])
def test_fold_single_gpu_matches_oracle_and_two_pass_form(shape):
    from ssi_b200 import fold as sfold
    from ssi_b200.step import PathStep

    inp = _inputs(**shape)
    dev = torch.device("cuda", 0)
    args = [inp[k].to(dev) for k in ("tokens", "labels", "X", "dH0")]
    folded = PathStep(inp["W"].to(dev), fold_scale_grads=True)
    plain = PathStep(inp["W"].to(dev))
    runs = []
    for _ in range(2):  # same buffers twice: the partials of the first step must not leak into the second
        folded.zero_grad()
        folded(*args)
        sq = folded.fold.sqnorm()
        torch.cuda.synchronize()
        runs.append((folded.weight.grad.clone(), sq.item()))
    assert torch.equal(runs[0][0], runs[1][0]) and runs[0][1] == runs[1][1]  # deterministic
    scaler = _check_against_oracle(inp, runs[0][0].cpu(), runs[0][1])
    # the two-pass form: unfolded step, then the reference's scale_grads on the bf16 gradient
    plain(*args)
    two_pass = plain.weight.grad.clone()
    two_pass *= torch.tensor(scaler, device=dev)
    torch.cuda.synchronize()
    a, b = runs[0][0].float(), two_pass.float()
    assert_grad_close(a.cpu(), b.cpu(), "folded vs two-pass gradient")
    # rows the embedding did not touch: one rounding against two, never more than one bf16 ulp apart; the touched rows
    # round their two parts at different points (covered by the tolerance above and by the oracle comparison)
    untouched = torch.ones(a.size(0), dtype=torch.bool, device=dev)
    ids = args[0].reshape(-1)
    untouched[ids[(ids >= 0) & (ids < a.size(0))]] = False
    au, bu = a[untouched], b[untouched]
    ulp = torch.maximum(bu.abs(), torch.tensor(1e-30, device=dev)) * 2.0 ** -7
    assert ((au - bu).abs() <= ulp).all(), "untouched rows: folded and two-pass form differ by more than one bf16 ulp"
    # helpers: scale_grads skips the folded weight, grad_norm takes the folded sum of squares
    before = folded.weight.grad.clone()
    sfold.scale_grads(folded.model, torch.tensor(123.0), folded.loss_fn)
    assert torch.equal(folded.weight.grad, before)
    norm = sfold.grad_norm(folded.model, folded.loss_fn)
    assert abs(norm.item() - runs[0][1] ** 0.5) <= 1e-6 * runs[0][1] ** 0.5


def test_fold_on_the_last_micro_batch_of_an_accumulation_window():
    from ssi_b200 import ops
    from ssi_b200.step import PathStep

    inp = _inputs(B=4, S=128, V=4001, D=256, seed=31)
    dev = torch.device("cuda", 0)
    args = [inp[k].to(dev) for k in ("tokens", "labels", "X", "dH0")]
    n_iter = int((inp["labels"] != -100).sum())
    scaler = torch.tensor([1.0 / (2 * n_iter)], device=dev)
    step = PathStep(inp["W"].to(dev), fold_scale_grads=True)
    step.fold_every_step = False  # the trainer, not PathStep, decides when to fold
    step(*args)            # micro-batch 1 of the window: plain accumulation
    step.fold.arm(scaler)  # before the forward of the window's last micro-batch
    step(*args)
    sq = step.fold.sqnorm()
    torch.cuda.synchronize()
    _check_against_oracle(inp, step.weight.grad.cpu(), sq.item(), micro_batches=2)
    # against the library's own stand-alone pass on an unfolded window
    plain = PathStep(inp["W"].to(dev))
    plain(*args)
    plain(*args)
    ref_sq = ops.grad_scale_sqnorm(plain.weight.grad, scaler)
    torch.cuda.synchronize()
    assert abs(sq.item() - ref_sq.item()) <= 2e-3 * ref_sq.item()


def test_fold_through_the_peer_exchange_world1():
    from ssi_b200.step import PathStep

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=0, world_size=1)
    try:
        inp = _inputs(seed=41)
        dev = torch.device("cuda", 0)
        args = [inp[k].to(dev) for k in ("tokens", "labels", "X", "dH0")]
        step = PathStep(inp["W"].to(dev), process_group=dist.group.WORLD, exchange="peer", fold_scale_grads=True)
        outs = []
        for _ in range(2):
            step.zero_grad()
            step(*args)
            sq = step.fold.sqnorm()
            torch.cuda.synchronize()
            outs.append((step.weight.grad.clone(), sq.item()))
        assert torch.equal(outs[0][0], outs[1][0]) and outs[0][1] == outs[1][1]
        _check_against_oracle(inp, outs[0][0].cpu(), outs[0][1])
        step.close()
    finally:
        dist.destroy_process_group()


def test_fold_needs_the_library_kernels():
    from ssi_b200 import install

    class M(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.tok_embeddings = torch.nn.Embedding(100, 64, dtype=torch.bfloat16, device="cuda")
            self.output = None

    with pytest.raises(ValueError):
        install(M(), mode="recompute", fold_scale_grads=True)
    with pytest.raises(ValueError):
        install(M(), replace_embedding=False, fold_scale_grads=True)
