"""GPU tests of the reference-facing surface: the loss module called like ssi/loss.py:20 calls it, the model
hooks, and the whole micro-batch (PathStep) against the golden fixtures produced by the REFERENCE's own modules
(tests/golden/, oracle/gen_golden.py) and against the CPU oracle.  Tolerances as in test_gpu_parity.py."""
import glob
import os

import pytest
import torch
from torch import nn

from oracle import reference_path as oracle
from test_gpu_parity import DW_ELEM_ROUNDINGS, LOSS_RTOL, assert_grad_close

pytestmark = pytest.mark.gpu

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.pt")))


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_path_step_matches_reference_fixture(path):
    from ssi_b200.step import PathStep

    fx = torch.load(path)
    inp, ref = fx["inputs"], fx["fp32"]
    dev = torch.device("cuda")
    step = PathStep(inp["W"].to(dev), num_output_chunks=fx["chunks"])
    loss_batch, dX, counts = step(inp["tokens"].to(dev), inp["labels"].to(dev), inp["X"].to(dev), inp["dH0"].to(dev))
    torch.cuda.synchronize()
    want = ref["loss"].item() * ref["num_tokens_iter"]
    assert counts[0].item() == ref["num_tokens_iter"]
    assert abs(loss_batch.item() - want) <= LOSS_RTOL * abs(want)
    assert_grad_close(dX.reshape(ref["dH"].shape), ref["dH"], "dH")
    assert_grad_close(step.weight.grad, ref["dW"], "dW(head+embedding)", elem_roundings=DW_ELEM_ROUNDINGS)


def test_loss_module_accepts_chunk_list_and_flat_inputs():
    import ssi_b200
    from ssi_b200 import synth

    inp = synth.make_inputs(synth.CONFIGS["c2"], batch=3, seq=50, vocab=777, hidden=128, seed=9)
    dev = torch.device("cuda")
    emb = nn.Embedding(777, 128, dtype=torch.bfloat16, device=dev)
    with torch.no_grad():
        emb.weight.copy_(inp["W"])
    loss_fn = ssi_b200.FusedLinearCrossEntropyLoss(emb)
    shifted = oracle.shift_labels(inp["labels"])
    want = oracle.lmhead_ce(inp["X"].float(), inp["W"].float(), shifted).item()
    h = inp["X"].to(dev)
    # (1) list of chunks + [b, s] labels, as compute_loss passes them when the model output is a list
    chunks = list(h.chunk(8, dim=1))  # s=50 -> 8 chunks of 7,7,7,7,7,7,7,1
    l1 = loss_fn(chunks, shifted.to(dev))
    # (2) flattened [b*s, d] + [b*s] (ssi/loss.py:17-19)
    l2 = loss_fn(h.reshape(-1, 128), shifted.reshape(-1).to(dev))
    # (3) forward only under inference_mode (ssi/eval.py:26-33)
    with torch.inference_mode():
        l3 = loss_fn(chunks, shifted.to(dev))
    for l in (l1, l2, l3):
        assert l.dtype == torch.float32 and l.dim() == 0
        assert abs(l.item() - want) <= LOSS_RTOL * abs(want)
    assert l1.item() == l2.item() == l3.item()  # deterministic, layout independent


def test_installed_model_trains_like_reference_flow():
    """Mirror of Trainer._setup_loss/_train_step (ssi/trainer.py:299-304,385-395) on a tiny tied model."""
    import ssi_b200
    from ssi_b200 import synth

    B, S, V, D = 2, 64, 515, 64
    inp = synth.make_inputs(synth.CONFIGS["c3"], batch=B, seq=S, vocab=V, hidden=D, seed=4)
    dev = torch.device("cuda")

    class TinyDecoder(nn.Module):
        def __init__(self):
            super().__init__()
            self.tok_embeddings = nn.Embedding(V, D, dtype=torch.bfloat16)
            self.output = None
            self.num_output_chunks = 0
            self.mix = nn.Parameter(torch.tensor(0.5, dtype=torch.bfloat16))

        def set_num_output_chunks(self, n):
            self.num_output_chunks = n

        def forward(self, tokens, **kw):
            h = self.tok_embeddings(tokens) * self.mix + extra
            return [self.output(c) for c in h.chunk(self.num_output_chunks, dim=1)]

    extra = inp["X"].to(dev)
    model = TinyDecoder().to(dev)
    with torch.no_grad():
        model.tok_embeddings.weight.copy_(inp["W"])
    loss_fn = ssi_b200.install(model)
    tokens, labels = inp["tokens"].to(dev), inp["labels"].to(dev)
    n_iter = int((labels != loss_fn.ignore_index).sum().item())
    logits = model(tokens=tokens)
    shifted = torch.hstack((labels[..., 1:], torch.full_like(labels[..., -1:], loss_fn.ignore_index)))
    loss = loss_fn(logits, shifted) * n_iter
    loss.backward()
    torch.cuda.synchronize()

    # oracle: same tiny model with the reference modules' arithmetic, fp32 on CPU
    Wf = inp["W"].float().clone().requires_grad_(True)
    mix = torch.tensor(0.5, requires_grad=True)
    h = torch.nn.functional.embedding(inp["tokens"], Wf) * mix + inp["X"].float()
    ref = oracle.lmhead_ce(h, Wf, oracle.shift_labels(inp["labels"])) * n_iter
    ref.backward()
    assert abs(loss.item() - ref.item()) <= 2e-3 * abs(ref.item())  # h itself is rounded to bf16 in the model
    assert_grad_close(model.tok_embeddings.weight.grad, Wf.grad, "tied dW", elem_roundings=DW_ELEM_ROUNDINGS)
    assert abs(model.mix.grad.float().item() - mix.grad.item()) <= 2e-2 * abs(mix.grad.item()) + 1e-3


def test_gradient_accumulation_and_determinism():
    from ssi_b200 import synth
    from ssi_b200.step import PathStep

    dev = torch.device("cuda")
    V, D = 2050, 128
    a = synth.make_inputs(synth.CONFIGS["c3"], batch=2, seq=96, vocab=V, hidden=D, seed=1)
    b = synth.make_inputs(synth.CONFIGS["c2"], batch=2, seq=96, vocab=V, hidden=D, seed=2)
    W = a["W"]

    def run():
        step = PathStep(W.to(dev))
        outs = []
        for mb in (a, b):  # two micro-batches accumulate into the same .grad (grad accumulation, conf/training.yaml:11)
            outs.append(step(mb["tokens"].to(dev), mb["labels"].to(dev), mb["X"].to(dev), mb["dH0"].to(dev)))
        torch.cuda.synchronize()
        return step.weight.grad.clone(), outs

    g1, o1 = run()
    g2, o2 = run()
    assert torch.equal(g1, g2)  # bit-identical run to run (tests/test_resume_equivalence.py relies on this)
    assert all(torch.equal(x[1], y[1]) and x[0].item() == y[0].item() for x, y in zip(o1, o2))
    ref = torch.zeros(V, D)
    for mb in (a, b):
        _, _, _, dW = oracle.train_step_path(mb["tokens"], mb["labels"], mb["X"], mb["dH0"], W)
        ref += dW
    assert_grad_close(g1, ref, "accumulated dW", elem_roundings=DW_ELEM_ROUNDINGS)


@pytest.mark.parametrize("ctxmgr", [torch.inference_mode, torch.no_grad])
def test_eval_path_is_forward_only(ctxmgr):
    """ssi/eval.py:27-33 calls the loss under inference mode: only the online-LSE kernel may run there (no logits
    block, no dW / dX contraction), and the value equals the training-mode loss."""
    from ssi_b200 import _lib
    from ssi_b200.step import PathStep

    fx = torch.load(GOLDEN[0])
    inp = fx["inputs"]
    dev = torch.device("cuda")
    step = PathStep(inp["W"].to(dev))
    X, labels = inp["X"].to(dev), inp["labels"].to(dev)
    _lib.profile_begin()
    with ctxmgr():
        shifted = torch.hstack((labels[..., 1:], torch.full_like(labels[..., :1], -100)))  # ssi/loss.py:16
        loss = step.loss_fn(X, shifted)
    cls = _lib.profile_end()
    assert cls["lse_fwd"][1] >= 1
    assert cls["dW"][1] == 0 and cls["dX"][1] == 0 and cls["softmax_grad"][1] == 0
    want = fx["fp32"]["loss"].item()
    assert abs(loss.item() - want) <= LOSS_RTOL * abs(want)
