"""Parity at BASELINE.json's FULL sizes (configs c1-c4, and c5's per-rank shares at 8, 4, 2 and 1 GPUs) through the
C ABI.

The oracle's arithmetic (oracle/reference_path.py: the reference's own torch ops) is too slow on the CPU at these
sizes, so it is evaluated here on the GPU tensors: the whole-batch loss by the oracle's chunked CE (forward only,
bf16 logits like the reference), and the gradients on SAMPLES that the oracle's formula gives in closed form —
256 token rows of dX and 256 vocabulary rows of dW, recomputed in fp32 from the same inputs.  Size-independent
properties close the rest: ignored rows have exactly zero dX, doubling the upstream gradient doubles every
gradient element exactly (a power of two commutes with every bf16 rounding on the path), and a second run is
bit-identical.  Tolerances as in test_gpu_parity.py."""
import pytest
import torch

from oracle import reference_path as oracle
from test_gpu_parity import DW_ELEM_ROUNDINGS, LOSS_RTOL, assert_grad_close

pytestmark = pytest.mark.gpu

# (config, batch override): c5 at 8 / 16 / 32 / 64 sequences = the per-rank shapes bench.py times at 8 / 4 / 2 / 1 GPUs
# (1 / 2 / 4 / 8 row chunks at V = 133,258)
FULL = [("c1", None), ("c2", None), ("c3", None), ("c4", None), ("c5", 8), ("c5", 16), ("c5", 32), ("c5", 64)]


def _run(X, W, shifted, upstream):
    from ssi_b200 import ops

    N, D = X.shape
    h = torch.tensor([upstream], dtype=torch.float32, device=X.device)
    loss_out, lse, n_valid, dX, dW_acc = ops.lmhead_ce_fwd_grad(X, W, shifted, h)
    dX, dW = ops.lmhead_ce_apply_grad(dX, dW_acc, h, h)
    return loss_out, lse, n_valid, dX, dW


@pytest.mark.parametrize("name,batch", FULL, ids=[n if b is None else f"{n}-{b}seq" for n, b in FULL])
def test_full_size_parity(name, batch):
    from ssi_b200 import synth

    cfg = synth.CONFIGS[name]
    B, S, V, D = batch or cfg.batch, cfg.seq, cfg.vocab, synth.HIDDEN
    dev = torch.device("cuda")
    g = torch.Generator().manual_seed(synth.SEED + 1)
    tokens, labels = synth.make_tokens_labels(cfg, g, batch=B)
    gd = torch.Generator(device=dev).manual_seed(synth.SEED)
    W = (torch.randn(V, D, generator=gd, device=dev) * 0.02).to(torch.bfloat16)
    X = torch.randn(B * S, D, generator=gd, device=dev).to(torch.bfloat16)
    shifted = oracle.shift_labels(labels).reshape(-1).to(dev)
    valid = shifted != -100
    n_valid = int(valid.sum())
    upstream = float((labels != -100).sum())  # ssi/trainer.py:391-393

    loss_out, lse, nv, dX, dW = _run(X, W, shifted, upstream)
    torch.cuda.synchronize()
    assert nv.item() == n_valid

    # ---- whole-batch loss: the oracle's chunked CE on bf16 logits (the reference dtype), forward only
    with torch.no_grad():
        chunks = max(8, (B * S) // 4096)
        ref_loss = oracle.lmhead_ce(X.reshape(1, B * S, D), W, shifted.reshape(1, -1), num_output_chunks=chunks)
    assert abs(loss_out[1].item() - ref_loss.item()) <= LOSS_RTOL * abs(ref_loss.item())

    scale = upstream / n_valid
    gs = torch.Generator().manual_seed(7)
    # ---- 256 sampled token rows (valid and ignored): lse and dX from the oracle's formula in fp32
    rows = torch.randperm(B * S, generator=gs)[:256].to(dev)
    with torch.no_grad():
        logits = (X[rows].float() @ W.float().t()).to(torch.bfloat16).float()
        ref_lse = torch.logsumexp(logits, dim=-1)
        rv = valid[rows]
        assert torch.allclose(lse[rows][rv], ref_lse[rv], rtol=1e-4, atol=2e-3)
        G = torch.softmax(logits, dim=-1)
        G[torch.arange(256, device=dev)[rv], shifted[rows][rv]] -= 1.0
        G = (G * scale * rv[:, None]).to(torch.bfloat16).float()  # autograd hands bf16 dlogits to the matmuls
        ref_dX = G @ W.float()
    assert_grad_close(dX[rows], ref_dX, "dX (sampled rows)")
    assert (dX[~valid] == 0).all()

    # ---- 256 sampled vocabulary rows (label ids included): dW from the oracle's formula in fp32
    lab_ids = shifted[valid][torch.randperm(n_valid, generator=gs)[:128].to(dev)]
    vs = torch.unique(torch.cat([lab_ids, torch.randint(0, V, (128,), generator=gs).to(dev)]))
    with torch.no_grad():
        Xv = X[valid]
        lg = (Xv.float() @ W[vs].float().t()).to(torch.bfloat16).float()  # [n_valid, |vs|]
        Gv = torch.exp(lg - lse[valid][:, None])
        Gv -= (shifted[valid][:, None] == vs[None, :]).float()
        Gv = (Gv * scale).to(torch.bfloat16).float()
        ref_dW = Gv.t() @ Xv.float()
    assert_grad_close(dW[vs], ref_dW, "dW (sampled vocabulary rows)", elem_roundings=DW_ELEM_ROUNDINGS)

    # ---- properties over the WHOLE result
    loss2, lse2, _, dX2, dW2 = _run(X, W, shifted, 2 * upstream)
    torch.cuda.synchronize()
    assert torch.equal(loss2, loss_out) and torch.equal(lse2, lse)
    assert torch.equal(dX2.float(), 2 * dX.float()), "doubling the upstream gradient must double dX exactly"
    assert torch.equal(dW2.float(), 2 * dW.float()), "doubling the upstream gradient must double dW exactly"
    _, _, _, dX3, dW3 = _run(X, W, shifted, upstream)
    torch.cuda.synchronize()
    assert torch.equal(dX3, dX) and torch.equal(dW3, dW), "run-to-run determinism"
