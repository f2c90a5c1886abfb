"""GPU parity of SINGLE-PASS mode (ssi_b200_lmhead_ce_fwd_grad / _apply_grad through the C ABI): forward also
forms dX and the fp32 dW sum from a bounded bf16 logits block, each logits tile computed once.

Same oracle and tolerances as test_gpu_parity.py: loss rel. err <= 1e-3; dX, dW rel-Frobenius <= 4e-3 and per
element |err| <= 2^-7 |ref| + 2^-8 max|ref| against the fp32 oracle (dW: x DW_ELEM_ROUNDINGS, see there).  The per-row logsumexp is taken on
bf16-rounded logits (the reference's logits dtype), so it is compared with the oracle run in bf16 as well.
"""
import pytest
import torch

from oracle import reference_path as oracle
from test_gpu_parity import CASES, DW_ELEM_ROUNDINGS, FRO_TOL, LOSS_RTOL, _case, assert_grad_close

pytestmark = pytest.mark.gpu


@pytest.fixture
def small_blocks():
    """Force several row chunks: cap the logits block at 1 MiB for the duration of a test."""
    from ssi_b200 import _lib

    _lib.set_option("g_cap_mib", 1)
    yield
    _lib.set_option("g_cap_mib", 4608)


def _run(inp, D, upstream, hint=None):
    from ssi_b200 import ops

    X = inp["X"].reshape(-1, D).cuda()
    W = inp["W"].cuda()
    lab = inp["shifted"].reshape(-1).cuda()
    h = torch.tensor([upstream if hint is None else hint], dtype=torch.float32, device="cuda")
    loss_out, lse, n_valid, dX, dW_acc = ops.lmhead_ce_fwd_grad(X, W, lab, h)
    g = torch.tensor([upstream], dtype=torch.float32, device="cuda")
    dX, dW = ops.lmhead_ce_apply_grad(dX, dW_acc, g, h)
    torch.cuda.synchronize()
    return loss_out, lse, n_valid, dX, dW, dW_acc


def _check(inp, D, out, upstream):
    loss_out, lse, n_valid, dX, dW, _ = out
    ref_loss = oracle.lmhead_ce(inp["X"].float(), inp["W"].float(), inp["shifted"])
    assert n_valid.item() == oracle.count_valid(inp["shifted"])
    assert abs(loss_out[1].item() - ref_loss.item()) <= LOSS_RTOL * abs(ref_loss.item())
    # lse on bf16-rounded logits, as the reference computes it (tied_linear.py:25 -> ce_chunked_output_loss.py:41)
    logits_bf16 = (inp["X"].reshape(-1, D).float() @ inp["W"].float().t()).to(torch.bfloat16).float()
    ref_lse = torch.logsumexp(logits_bf16, dim=-1)
    valid = inp["shifted"].reshape(-1) != -100
    assert torch.allclose(lse.cpu()[valid], ref_lse[valid], rtol=1e-4, atol=2e-3)
    assert (lse.cpu()[~valid] == 0).all()
    _, rdX, rdW = oracle.lmhead_ce_fwd_bwd(inp["X"], inp["W"], inp["shifted"], upstream)
    assert_grad_close(dX, rdX.reshape(-1, D), "dX")
    assert_grad_close(dW, rdW, "dW", elem_roundings=DW_ELEM_ROUNDINGS)
    assert (dX.float().cpu()[~valid] == 0).all()


@pytest.mark.parametrize("cfg,B,S,V,D", CASES)
def test_single_pass_matches_oracle(cfg, B, S, V, D):
    inp = _case(cfg, B, S, V, D, seed=1)
    upstream = float(oracle.count_valid(inp["labels"]))
    _check(inp, D, _run(inp, D, upstream), upstream)


@pytest.mark.parametrize("cfg,B,S,V,D", [("c3", 4, 300, 4098, 256), ("c2", 8, 256, 5002, 128), ("c4", 3, 700, 9000, 64)])
def test_single_pass_many_row_chunks(small_blocks, cfg, B, S, V, D):
    from ssi_b200 import _lib

    inp = _case(cfg, B, S, V, D, seed=2)
    upstream = float(oracle.count_valid(inp["labels"]))
    out = _run(inp, D, upstream)
    _check(inp, D, out, upstream)
    # same inputs with one big block: the chunking changes the fp32 summation order of dW and -- through the split
    # schedule of the logits contraction -- the last bits of the row sums, hence a few bf16 roundings of the scaled rows
    _lib.set_option("g_cap_mib", 4608)
    one = _run(inp, D, upstream)
    assert_grad_close(out[3], one[3], "dX(chunked vs one block)")
    fro = (out[5] - one[5]).norm().item() / one[5].norm().item()
    assert fro <= 1e-3, f"dW_acc(chunked vs one block): relative Frobenius difference {fro:.3e}"


def test_wrong_hint_is_rescaled_and_accumulate():
    cfg, B, S, V, D = CASES[2]
    inp = _case(cfg, B, S, V, D, seed=3)
    upstream = float(oracle.count_valid(inp["labels"]))
    from ssi_b200 import ops

    out = _run(inp, D, upstream, hint=1.0)  # hint 1.0, real upstream = token count -> ratio path
    _check(inp, D, out, upstream)
    # accumulate into an existing gradient buffer
    _, _, _, dX, dW, dW_acc = out
    _, _, rdW = oracle.lmhead_ce_fwd_bwd(inp["X"], inp["W"], inp["shifted"], upstream)
    base = (torch.randn(V, D) * rdW.abs().max()).to(torch.bfloat16)
    dW2 = base.clone().cuda()
    g = torch.tensor([upstream], dtype=torch.float32, device="cuda")
    ops.lmhead_ce_apply_grad(dX.clone(), dW_acc, g, torch.ones(1, device="cuda"), dW=dW2, accumulate_dW=True)
    torch.cuda.synchronize()
    assert_grad_close(dW2, rdW + base.float(), "dW(accumulate)", elem_roundings=DW_ELEM_ROUNDINGS)


def test_single_pass_deterministic_and_all_ignored():
    from ssi_b200 import ops

    cfg, B, S, V, D = CASES[3]
    inp = _case(cfg, B, S, V, D, seed=4)
    a = _run(inp, D, 17.0)
    b = _run(inp, D, 17.0)
    for x, y in zip(a, b):
        assert torch.equal(x, y)
    X = torch.randn(128, 64).to(torch.bfloat16).cuda()
    W = torch.randn(512, 64).to(torch.bfloat16).cuda()
    lab = torch.full((128,), -100, dtype=torch.int64).cuda()
    one = torch.ones(1, device="cuda")
    loss_out, lse, n_valid, dX, dW_acc = ops.lmhead_ce_fwd_grad(X, W, lab, one)
    assert n_valid.item() == 0 and torch.isnan(loss_out[1]).item()
    dX, dW = ops.lmhead_ce_apply_grad(dX, dW_acc, one, one)
    assert dX.float().abs().max().item() == 0 and dW.float().abs().max().item() == 0


def test_module_modes_agree():
    """FusedLinearCrossEntropyLoss in single_pass and recompute mode: same loss (to bf16-logit rounding) and
    gradients within the oracle tolerance of each other, through autograd with the trainer's multiplier."""
    from torch import nn

    from ssi_b200.loss import FusedLinearCrossEntropyLoss

    cfg, B, S, V, D = CASES[5]
    inp = _case(cfg, B, S, V, D, seed=5)
    res = {}
    for mode in ("single_pass", "recompute"):
        emb = nn.Embedding(V, D, dtype=torch.bfloat16, device="cuda")
        with torch.no_grad():
            emb.weight.copy_(inp["W"].cuda())
        loss_fn = FusedLinearCrossEntropyLoss(emb, mode=mode)
        h = inp["X"].cuda().requires_grad_(True)
        n_iter = torch.tensor(float(oracle.count_valid(inp["labels"])), device="cuda")
        loss_fn.grad_hint = n_iter.reshape(1)
        loss = loss_fn([c for c in h.chunk(8, dim=1)], inp["shifted"].cuda())
        (loss * n_iter).backward()
        torch.cuda.synchronize()
        res[mode] = (loss.item(), h.grad.clone(), emb.weight.grad.clone())
    assert abs(res["single_pass"][0] - res["recompute"][0]) <= 1e-3 * abs(res["recompute"][0])
    # two of our own results, each within the oracle tolerance: twice the bound between them
    assert_grad_close(res["single_pass"][1], res["recompute"][1], "dX(single vs recompute)", roundings=2.0)
    assert_grad_close(res["single_pass"][2], res["recompute"][2], "dW(single vs recompute)", roundings=2.0)


@pytest.mark.parametrize("exp_stash", [1, 0])
def test_large_logits_and_a_spike_far_above_the_first_vocabulary_tile(exp_stash):
    """The exp-stash form references every row's exponentials to the max of the FIRST vocabulary tile (a pilot sweep),
    not to the row's true max: logits 8x larger than usual and one logit ~50 nats above everything in the pilot tile
    must still give the right loss and gradients (bf16 has the exponent range of fp32).  With logits this large the
    reference's rounding of the logits to bf16 (tied_linear.py:25) moves the probabilities by several per cent, so the
    oracle here is fp32 arithmetic ON the bf16-rounded logits -- what the reference's fp32 cross-entropy sees."""
    from ssi_b200 import _lib

    g = torch.Generator().manual_seed(77)
    N, V, D = 384, 1500, 128
    X = (torch.randn(N, D, generator=g) * 2.0).to(torch.bfloat16)
    W = (torch.randn(V, D, generator=g) * (4.0 * D ** -0.5)).to(torch.bfloat16)  # logits ~ N(0, 8^2)
    W[900] = (X[5].float() * (75.0 / X[5].float().pow(2).sum())).to(torch.bfloat16)  # logit[5, 900] ~ 75
    labels = torch.randint(0, V, (N,), generator=g)
    labels[5] = 900
    labels[6] = 3
    inp = {"X": X.reshape(1, N, D), "W": W, "shifted": labels.reshape(1, N)}
    upstream = 100.0
    _lib.set_option("exp_stash", exp_stash)
    try:
        loss_out, lse, n_valid, dX, dW, _ = _run(inp, D, upstream)
    finally:
        _lib.set_option("exp_stash", 1)
    logits = (X.float() @ W.float().t()).to(torch.bfloat16).float()
    assert logits[5, 900] - logits[5, :256].max() > 45.0  # the spike is far above the pilot tile
    ref_lse = torch.logsumexp(logits, dim=-1)
    G = (torch.softmax(logits, dim=-1) - torch.nn.functional.one_hot(labels, V).float()) * (upstream / N)
    ref_loss = (ref_lse - logits[torch.arange(N), labels]).mean()
    assert abs(loss_out[1].item() - ref_loss.item()) <= LOSS_RTOL * abs(ref_loss.item())
    assert torch.allclose(lse.cpu(), ref_lse, rtol=1e-4, atol=2e-3)
    assert torch.isfinite(dX.float()).all() and torch.isfinite(dW.float()).all()
    assert_grad_close(dX, G @ W.float(), "dX")
    assert_grad_close(dW, G.t() @ X.float(), "dW", elem_roundings=DW_ELEM_ROUNDINGS)


def _confident_case(target_p, seed, B=2, S=300, V=5002, D=256):
    """Rows whose hidden state points along their label's embedding, scaled so that softmax[label] ~ target_p."""
    import math

    g = torch.Generator().manual_seed(seed)
    W = (torch.randn(V, D, generator=g) * 0.06).to(torch.bfloat16)
    labels = torch.randint(0, V, (B * S,), generator=g)
    t = math.log(V) + math.log(target_p / (1 - target_p))
    wl = W[labels].float()
    X = (wl * (t / wl.pow(2).sum(-1, keepdim=True))).to(torch.bfloat16)
    ignored = torch.rand(B * S, generator=g) < 0.2
    labels[ignored] = -100
    return {"X": X.reshape(B, S, D), "W": W, "shifted": labels.reshape(B, S)}


@pytest.mark.parametrize("target_p", [0.9, 0.99, 0.998])
def test_confident_rows_keep_the_reference_accuracy(target_p):
    """with a peaked softmax the label column's gradient g (p_label - 1) nearly cancels; the
    exp-stash form used to round p_label to bf16 inside E BEFORE subtracting the onehot, leaving an absolute error of
    2^-9 g |W| on a signal of g (1 - p_label) |W| (31 % at p_label = 0.998).  The label column is now handled in fp32
    (1 - p_label = (L - e_label) / L).  Logits of ~10-15 nats move by several per cent of a nat when rounded to bf16
    (tied_linear.py:25), so -- as in the spike test above -- the oracle is fp32 arithmetic ON the bf16-rounded logits,
    what the reference's fp32 cross-entropy sees.  Bounds: the usual ones overall, and per row at most 4e-3 or the error
    of the reference's own arithmetic (dlogits rounded to bf16 before the contractions, bf16 results)."""
    inp = _confident_case(target_p, seed=int(target_p * 1000))
    V, D = inp["W"].shape
    X, W = inp["X"].reshape(-1, D).float(), inp["W"].float()
    labels = inp["shifted"].reshape(-1)
    valid = labels != -100
    nv = int(valid.sum())
    upstream = float(nv)
    logits = (X @ W.t()).to(torch.bfloat16).float()
    P = torch.softmax(logits, -1)
    p_lab = P[valid][torch.arange(nv), labels[valid]]
    assert abs(p_lab.median().item() - target_p) < 0.5 * (1 - target_p) + 0.05, p_lab.median().item()
    G = torch.zeros_like(P)
    G[valid] = (P[valid] - torch.nn.functional.one_hot(labels[valid], V).float()) * (upstream / nv)
    rdX, rdW = G @ W, G.t() @ X
    Gb = G.to(torch.bfloat16).float()  # the reference's autograd hands bf16 dlogits to the two contractions
    bdX, bdW = (Gb @ W).to(torch.bfloat16).float(), (Gb.t() @ X).to(torch.bfloat16).float()
    loss_out, lse, n_valid, dX, dW, _ = _run(inp, D, upstream)
    assert torch.isfinite(dX.float()).all() and torch.isfinite(dW.float()).all()

    def row_err(got, ref):
        got, ref = got.float().cpu(), ref.float().cpu()
        return (got - ref).norm(dim=-1) / ref.norm(dim=-1).clamp_min(1e-30)

    ours, theirs = row_err(dX.cpu()[valid], rdX[valid]), row_err(bdX[valid], rdX[valid])
    touched = torch.unique(labels[valid])  # the label rows of dW carry the cancelling term
    ours_w, theirs_w = row_err(dW.cpu()[touched], rdW[touched]), row_err(bdW[touched], rdW[touched])
    print(f"p_label~{target_p}: dX rows ours {ours.median().item():.2e}/{ours.max().item():.2e} ref-bf16 "
          f"{theirs.median().item():.2e}/{theirs.max().item():.2e}; dW label rows ours "
          f"{ours_w.median().item():.2e}/{ours_w.max().item():.2e} ref-bf16 "
          f"{theirs_w.median().item():.2e}/{theirs_w.max().item():.2e}")
    assert_grad_close(dX, rdX, "dX")
    assert_grad_close(dW, rdW, "dW", elem_roundings=DW_ELEM_ROUNDINGS)
    assert ours.max().item() <= max(FRO_TOL, theirs.max().item()), (
        f"dX rows: worst relative error {ours.max().item():.3e} (median {ours.median().item():.3e}); the reference's "
        f"bf16 arithmetic: worst {theirs.max().item():.3e} (median {theirs.median().item():.3e})")
    assert ours_w.max().item() <= max(FRO_TOL, theirs_w.max().item()), (
        f"dW label rows: worst relative error {ours_w.max().item():.3e} (median {ours_w.median().item():.3e}); the "
        f"reference's bf16 arithmetic: worst {theirs_w.max().item():.3e} (median {theirs_w.median().item():.3e})")


def test_exp_stash_overflow_is_flagged_not_silent():
    """A logit more than ~88 nats above everything in the row's first 256 vocabulary columns overflows the stashed
    exponentials.  The call must say so (bit 1 of the status word, non-finite loss and gradients) instead of handing
    back plausible numbers; the logits-stash form (exp_stash = 0) has no such limit and stays exact."""
    from ssi_b200 import _lib, ops

    g = torch.Generator().manual_seed(5)
    N, V, D = 256, 1500, 128
    X = torch.randn(N, D, generator=g).to(torch.bfloat16)
    W = (torch.randn(V, D, generator=g) * D ** -0.5).to(torch.bfloat16)  # logits ~ N(0, 1)
    W[900] = (X[5].float() * (120.0 / X[5].float().pow(2).sum())).to(torch.bfloat16)  # logit[5, 900] ~ 120
    labels = torch.randint(0, V, (N,), generator=g)
    inp = {"X": X.reshape(1, N, D), "W": W, "shifted": labels.reshape(1, N)}
    logits = (X.float() @ W.float().t()).to(torch.bfloat16).float()
    assert logits[5, 900] - logits[5, :256].max() > 100.0
    loss_out, lse, n_valid, dX, dW, _ = _run(inp, D, 1.0)
    assert ops.single_pass_status(N, V, D) & 2, "overflow not flagged"
    assert not torch.isfinite(loss_out[1]).item() and not torch.isfinite(dX[5].float()).all()
    # an ordinary call on the same workspace clears the flag
    inp2 = dict(inp, W=(torch.randn(V, D, generator=g) * D ** -0.5).to(torch.bfloat16))
    _run(inp2, D, 1.0)
    assert ops.single_pass_status(N, V, D) & 2 == 0
    _lib.set_option("exp_stash", 0)
    try:
        loss_out, lse, n_valid, dX, dW, _ = _run(inp, D, 1.0)
    finally:
        _lib.set_option("exp_stash", 1)
    ref_lse = torch.logsumexp(logits, dim=-1)
    ref_loss = (ref_lse - logits[torch.arange(N), labels]).mean()
    assert abs(loss_out[1].item() - ref_loss.item()) <= LOSS_RTOL * abs(ref_loss.item())
    assert torch.isfinite(dX.float()).all() and torch.isfinite(dW.float()).all()
