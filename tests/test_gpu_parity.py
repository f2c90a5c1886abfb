"""GPU parity of the fused LM-head + CE path (called through the C ABI via ssi_b200.ops / the loss module)
against the CPU oracle (oracle/reference_path.py, fp32) on the same seeded inputs.

Tolerances: loss relative error <= 1e-3 (observed ~1e-6);
dX, dW: relative Frobenius error <= 4e-3 (2^-8) and per element |err| <= 2^-7 |ref| + 2^-8 max|ref|.
"""
import pytest
import torch

from oracle import reference_path as oracle

pytestmark = pytest.mark.gpu

LOSS_RTOL = 1e-3
FRO_TOL = 4e-3


# dW of single-pass mode contracts TWO bf16-rounded operands per token -- E = exp(logit - ref) and the row of X scaled
# by the softmax normalisation -- where the reference contracts one rounded operand (dlogits) with the exact X row.
# The rms error against the fp32 oracle stays inside the rel-Frobenius bound; the PER-ELEMENT bound, which the survey
# derived from the reference's own one-rounding bf16 path, is widened by this factor for dW only.
DW_ELEM_ROUNDINGS = 1.0  # per-element bound of the dW checks, in units of the bf16-output rounding bound


def assert_grad_close(got: torch.Tensor, ref: torch.Tensor, name: str, roundings: float = 1.0,
                      elem_roundings: float = 1.0):
    """``roundings`` > 1 widens both bounds for results that were ACCUMULATED in bf16 that many times (the NCCL
    baseline sums micro-batches and ranks in bf16, as the reference's autograd sums its chunks in bf16) or for two of
    our own results compared with each other; ``elem_roundings`` widens the per-element bound only."""
    got = got.float().cpu()
    ref = ref.float().cpu()
    fro = (got - ref).norm().item() / (ref.norm().item() + 1e-30)
    assert fro <= FRO_TOL * roundings, f"{name}: relative Frobenius error {fro:.3e} > {FRO_TOL * roundings}"
    bound = (ref.abs() * 2.0 ** -7 + ref.abs().max() * 2.0 ** -8) * roundings * elem_roundings
    over = (got - ref).abs() / bound
    bad = (over > 1).sum().item()
    assert bad == 0, (f"{name}: {bad} elements outside bf16 tolerance (worst {over.max().item():.2f} x the bound, "
                      f"fro {fro:.3e})")


def _case(cfg_name, batch, seq, vocab, hidden, seed=0):
    from ssi_b200 import synth

    cfg = synth.CONFIGS[cfg_name]
    inp = synth.make_inputs(cfg, batch=batch, seq=seq, vocab=vocab, hidden=hidden, seed=synth.SEED + seed)
    inp["shifted"] = oracle.shift_labels(inp["labels"])
    return inp


CASES = [
    # cfg, B, S, V, D
    ("c1", 1, 256, 1000, 64),       # single tile column, ragged V
    ("c3", 2, 128, 4098, 256),      # V = 2 mod 4 like every real vocabulary
    ("c2", 4, 192, 5002, 512),      # ~75% ignored rows, N not a multiple of 128
    ("c4", 2, 520, 9000, 128),      # packed rows with padding tail
    ("c3", 1, 100, 300, 64),        # N < one tile
    ("c1", 3, 384, 20010, 2048),    # real hidden size
]


@pytest.mark.parametrize("cfg,B,S,V,D", CASES)
def test_forward_loss_and_lse(cfg, B, S, V, D):
    from ssi_b200 import ops

    inp = _case(cfg, B, S, V, D)
    X = inp["X"].reshape(-1, D).cuda()
    W = inp["W"].cuda()
    lab = inp["shifted"].reshape(-1).cuda()
    loss_out, lse, n_valid = ops.lmhead_ce_fwd(X, W, lab)
    torch.cuda.synchronize()
    ref_loss = oracle.lmhead_ce(inp["X"].float(), inp["W"].float(), inp["shifted"])
    ref_lse = torch.logsumexp(inp["X"].reshape(-1, D).float() @ inp["W"].float().t(), dim=-1)
    assert n_valid.item() == oracle.count_valid(inp["shifted"])
    assert abs(loss_out[1].item() - ref_loss.item()) <= LOSS_RTOL * abs(ref_loss.item())
    valid = inp["shifted"].reshape(-1) != -100
    assert torch.allclose(lse.cpu()[valid], ref_lse[valid], rtol=1e-4, atol=1e-4)
    assert (lse.cpu()[~valid] == 0).all()  # ignored rows are compacted away and never reach the tensor cores


@pytest.mark.parametrize("cfg,B,S,V,D", CASES)
def test_backward_dx_dw(cfg, B, S, V, D):
    from ssi_b200 import ops

    inp = _case(cfg, B, S, V, D, seed=1)
    X = inp["X"].reshape(-1, D).cuda()
    W = inp["W"].cuda()
    lab = inp["shifted"].reshape(-1).cuda()
    upstream = float(oracle.count_valid(inp["labels"]))  # ssi/trainer.py:391-393
    loss_out, lse, n_valid = ops.lmhead_ce_fwd(X, W, lab)
    g = torch.tensor([upstream], dtype=torch.float32, device="cuda")
    dX, dW = ops.lmhead_ce_bwd(X, W, lab, lse, g, n_valid)
    torch.cuda.synchronize()
    _, rdX, rdW = oracle.lmhead_ce_fwd_bwd(inp["X"], inp["W"], inp["shifted"], upstream)
    assert_grad_close(dX, rdX.reshape(-1, D), "dX")
    assert_grad_close(dW, rdW, "dW")
    # accumulate into an existing gradient buffer
    base = (torch.randn(V, D) * rdW.abs().max()).to(torch.bfloat16)
    dW2 = base.clone().cuda()
    ops.lmhead_ce_bwd(X, W, lab, lse, g, n_valid, dW=dW2, accumulate_dW=True)
    torch.cuda.synchronize()
    assert_grad_close(dW2, rdW + base.float(), "dW(accumulate)")


@pytest.mark.parametrize("N,V,D,pattern", [
    (1000, 5002, 2048, "random"),
    (4096, 300, 256, "few_ids"),        # long runs crossing many warps
    (777, 9000, 64, "all_same"),
    (64, 100, 128, "sorted_unique"),
    (5000, 133258, 2048, "with_invalid"),
])
def test_embedding_grad_scatter(N, V, D, pattern):
    from ssi_b200 import ops

    g = torch.Generator().manual_seed(7)
    if pattern == "random" or pattern == "with_invalid":
        ids = torch.randint(0, V, (N,), generator=g)
    elif pattern == "few_ids":
        ids = torch.randint(0, 5, (N,), generator=g) * 50
    elif pattern == "all_same":
        ids = torch.full((N,), 1234, dtype=torch.int64)
    else:
        ids = torch.arange(N, dtype=torch.int64)
    dH = (torch.randn(N, D, generator=g) * 1e-3).to(torch.bfloat16)
    base = (torch.randn(V, D, generator=g) * 1e-3).to(torch.bfloat16)
    ids_dev = ids.clone()
    if pattern == "with_invalid":
        ids_dev[::97] = -1
        ids_dev[5::101] = V + 3
    keep = (ids_dev >= 0) & (ids_dev < V)
    ref = base.float() + oracle.embedding_grad(ids_dev[keep], dH[keep], V)
    dW = base.clone().cuda()
    ops.embedding_grad_scatter(dH.cuda(), ids_dev.cuda(), dW)
    torch.cuda.synchronize()
    assert_grad_close(dW, ref, "dW(embedding)")
    # deterministic: bit-identical on a second run
    dW_b = base.clone().cuda()
    ops.embedding_grad_scatter(dH.cuda(), ids_dev.cuda(), dW_b)
    torch.cuda.synchronize()
    assert torch.equal(dW, dW_b)


def test_label_shift_count():
    from ssi_b200 import ops

    g = torch.Generator().manual_seed(3)
    labels = torch.randint(0, 1000, (5, 77), generator=g)
    labels[torch.rand(5, 77, generator=g) < 0.3] = -100
    out, counts = ops.label_shift_count(labels.cuda())
    torch.cuda.synchronize()
    ref = oracle.shift_labels(labels)
    assert torch.equal(out.cpu(), ref)
    assert counts[0].item() == oracle.count_valid(labels)
    assert counts[1].item() == oracle.count_valid(ref)


def test_all_ignored_is_nan_and_zero_grads():
    from ssi_b200 import ops

    X = torch.randn(128, 64).to(torch.bfloat16).cuda()
    W = torch.randn(512, 64).to(torch.bfloat16).cuda()
    lab = torch.full((128,), -100, dtype=torch.int64).cuda()
    loss_out, lse, n_valid = ops.lmhead_ce_fwd(X, W, lab)
    assert n_valid.item() == 0 and torch.isnan(loss_out[1]).item()  # 0/0 as ce_chunked_output_loss.py:83
    dX, dW = ops.lmhead_ce_bwd(X, W, lab, lse, torch.ones(1, device="cuda"), n_valid)
    assert dX.float().abs().max().item() == 0 and dW.float().abs().max().item() == 0
