"""``compute_cross_entropy`` (tt/torchtune/modules/loss/ce_chunked_output_loss.py:35-43, the one
method ``training.compile_loss`` wraps, tt/torchtune/training/_compile.py:74-77, when ``cfg.compile`` is set,
ssi/trainer.py:301-302).  With the passthrough output installed the model hands over HIDDEN chunks, so the method
takes a hidden chunk, projects it inside the fused kernels and returns the SUM-reduced CE of the chunk -- compared
here with the oracle's ``ChunkedCE.compute_cross_entropy`` on the chunk's logits."""
import pytest
import torch
from torch import nn

from oracle import reference_path as oracle
from test_gpu_parity import DW_ELEM_ROUNDINGS, LOSS_RTOL, assert_grad_close

pytestmark = pytest.mark.gpu


def _setup(mode, B=3, S=64, V=1031, D=128, seed=21):
    import ssi_b200
    from ssi_b200 import synth

    inp = synth.make_inputs(synth.CONFIGS["c2"], batch=B, seq=S, vocab=V, hidden=D, seed=seed)
    dev = torch.device("cuda")
    emb = nn.Embedding(V, D, dtype=torch.bfloat16, device=dev)
    with torch.no_grad():
        emb.weight.copy_(inp["W"])
    loss_fn = ssi_b200.FusedLinearCrossEntropyLoss(emb, mode=mode)
    return inp, emb, loss_fn, dev


@pytest.mark.parametrize("mode", ["single_pass", "recompute"])
def test_compute_cross_entropy_matches_the_oracle_chunk_by_chunk(mode):
    inp, emb, loss_fn, dev = _setup(mode)
    shifted = oracle.shift_labels(inp["labels"])
    ref_fn = oracle.ChunkedCE(8)
    Wf = inp["W"].float().clone().requires_grad_(True)
    Xf = inp["X"].float().clone().requires_grad_(True)
    h = inp["X"].to(dev).requires_grad_(True)
    total, ref_total = 0.0, 0.0
    for hc, xc, lc in zip(h.chunk(8, dim=1), Xf.chunk(8, dim=1), shifted.chunk(8, dim=1)):
        got = loss_fn.compute_cross_entropy(hc, lc.to(dev))
        want = ref_fn.compute_cross_entropy(oracle.tied_linear(xc, Wf).reshape(-1, Wf.size(0)), lc.reshape(-1))
        assert got.dtype == torch.float32 and got.dim() == 0
        assert abs(got.item() - want.item()) <= LOSS_RTOL * abs(want.item()) + 1e-6
        total, ref_total = total + got, ref_total + want
    # the reference's forward is exactly this sum divided by the valid count (ce_chunked_output_loss.py:79-83)
    n = (shifted != -100).sum()
    (total / n.to(dev)).backward()
    (ref_total / n).backward()
    torch.cuda.synchronize()
    assert_grad_close(h.grad, Xf.grad, "dH through compute_cross_entropy")
    assert_grad_close(emb.weight.grad, Wf.grad, "dW through compute_cross_entropy", elem_roundings=DW_ELEM_ROUNDINGS)


@pytest.mark.parametrize("mode", ["single_pass", "recompute"])
def test_all_ignored_chunk_sums_to_zero_not_nan(mode):
    inp, emb, loss_fn, dev = _setup(mode, B=2, S=32)
    h = inp["X"].to(dev).requires_grad_(True)
    lab = torch.full((2, 32), -100, dtype=torch.int64, device=dev)
    got = loss_fn.compute_cross_entropy(h, lab)
    want = oracle.ChunkedCE().compute_cross_entropy(
        oracle.tied_linear(inp["X"].float(), inp["W"].float()).reshape(-1, inp["W"].size(0)), lab.cpu().reshape(-1))
    assert want.item() == 0.0 and got.item() == 0.0  # sum reduction: 0, where the mean would be 0/0
    got.backward()
    torch.cuda.synchronize()
    assert torch.count_nonzero(h.grad).item() == 0 and torch.count_nonzero(emb.weight.grad).item() == 0


def test_compiled_compute_cross_entropy_runs_forward_and_backward():
    """What ``compile_loss`` does (tt/torchtune/training/_compile.py:74-77): the method is replaced by its
    ``torch.compile``-d version; the fused call inside must neither crash dynamo nor change the numbers."""
    inp, emb, loss_fn, dev = _setup("single_pass")
    shifted = oracle.shift_labels(inp["labels"]).to(dev)
    h = inp["X"].to(dev)
    eager = loss_fn.compute_cross_entropy(h, shifted)
    import os

    backend = os.environ.get("TORCH_COMPILE_BACKEND", "inductor")  # the reference's choice (_compile.py:70)
    loss_fn.compute_cross_entropy = torch.compile(loss_fn.compute_cross_entropy, backend=backend)
    hc = h.clone().requires_grad_(True)
    out = loss_fn.compute_cross_entropy(hc, shifted)
    assert out.item() == eager.item()
    (out / (shifted != -100).sum()).backward()
    torch.cuda.synchronize()
    _, ref_dH, ref_dW = oracle.lmhead_ce_fwd_bwd(inp["X"], inp["W"], oracle.shift_labels(inp["labels"]))
    assert_grad_close(hc.grad, ref_dH, "dH, compiled compute_cross_entropy")
    assert_grad_close(emb.weight.grad, ref_dW, "dW, compiled compute_cross_entropy", elem_roundings=DW_ELEM_ROUNDINGS)
