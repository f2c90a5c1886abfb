"""GPU parity of the rows SURVEY.md §8(f) marks "next" (csrc/next_rows.cuh through the C ABI) against the oracle's
restatements of the reference call sites (oracle/reference_path.py):

  final RMSNorm fwd/bwd      tt/torchtune/modules/rms_norm.py:38-43        bf16 out: <= 1 bf16 ulp, >= 99.5 % exact
  token-type histogram       ssi/train_utils.py:150-165                    exact (integers)
  scale_grads + grad norm    tt/torchtune/training/_grad_scaler.py:25-31   scaled grad bit-exact, norm rel 1e-5
  embedding gather           tt/torchtune/modules/transformer.py:628       bit-exact
"""
import pytest
import torch

from oracle import reference_path as oracle

pytestmark = pytest.mark.gpu


def _ulp_close(got, ref, name, exact_frac=0.995):
    got, ref = got.float().cpu(), ref.float().cpu()
    ulp = torch.maximum(ref.abs(), torch.tensor(1e-30)) * 2.0 ** -7  # one bf16 ulp is at most 2^-7 |ref|
    assert ((got - ref).abs() <= ulp).all(), f"{name}: more than one bf16 ulp off"
    frac = (got == ref).float().mean().item()
    assert frac >= exact_frac, f"{name}: only {frac:.4f} of the elements are bit-exact"


@pytest.mark.parametrize("N,D", [(300, 2048), (129, 64), (1000, 520), (64, 4096), (17, 8192)])
def test_rmsnorm_fwd_bwd(N, D):
    from ssi_b200 import ops

    g = torch.Generator().manual_seed(N + D)
    x = (torch.randn(N, D, generator=g) * 3).to(torch.bfloat16)
    scale = (1 + 0.2 * torch.randn(D, generator=g)).to(torch.bfloat16)
    dy = (torch.randn(N, D, generator=g) * 1e-2).to(torch.bfloat16)
    ry, rdx, rds = oracle.rms_norm_fwd_bwd(x, scale, dy)
    y, rstd = ops.rmsnorm_fwd(x.cuda(), scale.cuda(), 1e-6)
    dx, dscale = ops.rmsnorm_bwd(dy.cuda(), x.cuda(), scale.cuda(), rstd)
    dx2, dscale2 = ops.rmsnorm_bwd(dy.cuda(), x.cuda(), scale.cuda(), rstd)
    torch.cuda.synchronize()
    _ulp_close(y, ry, "y")
    _ulp_close(dx, rdx, "dx", exact_frac=0.98)
    assert torch.allclose(dscale.cpu(), rds, rtol=1e-4, atol=1e-6 * max(1.0, rds.abs().max().item()))
    assert torch.equal(dx, dx2) and torch.equal(dscale, dscale2)  # deterministic


def test_rmsnorm_module_is_a_drop_in():
    from ssi_b200 import RMSNormB200

    torch.manual_seed(3)
    x = torch.randn(2, 50, 256).to(torch.bfloat16)
    scale = (1 + 0.1 * torch.randn(256)).to(torch.bfloat16)
    dy = torch.randn(2, 50, 256).to(torch.bfloat16)
    ry, rdx, rds = oracle.rms_norm_fwd_bwd(x.reshape(-1, 256), scale, dy.reshape(-1, 256))
    m = RMSNormB200(256, dtype=torch.bfloat16, device="cuda")
    with torch.no_grad():
        m.scale.copy_(scale)
    assert set(dict(m.named_parameters())) == {"scale"} and m.eps == 1e-6 and m.normalized_shape == (256,)
    xc = x.cuda().requires_grad_(True)
    y = m(xc)
    y.backward(dy.cuda())
    _ulp_close(y.reshape(-1, 256), ry, "y")
    _ulp_close(xc.grad.reshape(-1, 256), rdx, "dx", exact_frac=0.98)
    assert m.scale.grad.dtype == torch.bfloat16
    assert torch.allclose(m.scale.grad.float().cpu(), rds, rtol=2e-2, atol=1e-3)
    with torch.inference_mode():
        assert torch.equal(m(x.cuda()), y.detach())


def test_label_shift_count_types_matches_reference_counts():
    from ssi_b200 import ops, synth

    inp = synth.make_inputs(synth.CONFIGS["c2"], batch=5, seq=333, vocab=133258, hidden=64, seed=4)
    tokens, labels = inp["tokens"], inp["labels"]
    ranges = {"text": (0, 127999), "special_text": (128000, 128255), "dsu": (128256, 133255),
              "modality": (133256, 133257), "empty": (5, 4)}
    pad = int(tokens[0, -1])
    shifted, counts, types = ops.label_shift_count_types(labels.cuda(), tokens.cuda(), ranges, pad)
    s2, c2 = ops.label_shift_count(labels.cuda())
    torch.cuda.synchronize()
    assert torch.equal(shifted, s2) and torch.equal(counts, c2)
    assert torch.equal(shifted.cpu(), oracle.shift_labels(labels))
    ref = oracle.count_token_types(tokens, ranges, pad)
    got = dict(zip(list(ranges) + ["total"], types.tolist()))
    assert got == ref


@pytest.mark.parametrize("V,D", [(1000, 64), (5002, 2048)])
def test_grad_scale_sqnorm(V, D):
    from ssi_b200 import ops

    g = torch.Generator().manual_seed(V)
    grad = (torch.randn(V, D, generator=g) * 37).to(torch.bfloat16)
    scaler = torch.tensor(1 / 4099)  # ssi/trainer.py:404: torch.tensor(1 / num_tokens_step)
    ref_g, ref_sq = oracle.scale_grads_sqnorm(grad, scaler)
    gc = grad.cuda()
    sq = ops.grad_scale_sqnorm(gc, scaler.reshape(1).cuda())
    torch.cuda.synchronize()
    assert torch.equal(gc.cpu(), ref_g)
    assert abs(sq.item() - ref_sq.item()) <= 1e-5 * ref_sq.item()


def test_embedding_gather_and_tied_embedding_forward():
    from ssi_b200 import TiedEmbeddingB200, ops

    g = torch.Generator().manual_seed(8)
    W = torch.randn(3001, 512, generator=g).to(torch.bfloat16)
    ids = torch.randint(0, 3001, (3, 77), generator=g)
    out = ops.embedding_gather(W.cuda(), ids.cuda())
    torch.cuda.synchronize()
    assert torch.equal(out.cpu(), oracle.embedding_fwd(W, ids))
    emb = TiedEmbeddingB200(3001, 512, dtype=torch.bfloat16, device="cuda")
    with torch.no_grad():
        emb.weight.copy_(W)
    h = emb(ids.cuda())
    assert torch.equal(h.detach().cpu(), oracle.embedding_fwd(W, ids))
    dH = torch.randn(3, 77, 512, generator=g).to(torch.bfloat16)
    h.backward(dH.cuda())
    ref = oracle.embedding_grad(ids, dH, 3001)
    assert torch.allclose(emb.weight.grad.float().cpu(), ref, rtol=2 ** -7, atol=1e-2)


@pytest.mark.parametrize("mode", ["single_pass", "recompute"])
@pytest.mark.parametrize("B,S,V,D,cfg", [(3, 150, 3001, 256, "c2"), (2, 300, 5002, 2048, "c3")])
def test_final_norm_fused_into_the_path(B, S, V, D, cfg, mode):
    """install(fuse_norm=True) semantics: the loss takes PRE-norm hidden states plus the norm's scale and returns the
    loss and the gradients the reference gets from norm -> projection -> chunked CE (transformer.py:653, 499-502,
    ce_chunked_output_loss.py:45-83), including d(scale).  Oracle in fp32 on the same bf16 inputs."""
    import ssi_b200
    from ssi_b200 import synth
    from test_gpu_parity import DW_ELEM_ROUNDINGS, LOSS_RTOL, assert_grad_close

    inp = synth.make_inputs(synth.CONFIGS[cfg], batch=B, seq=S, vocab=V, hidden=D, seed=21)
    g = torch.Generator().manual_seed(5)
    h_pre = (inp["X"].float() * (1 + torch.rand(B, S, 1, generator=g) * 3)).to(torch.bfloat16)  # rows of varied rms
    scale = (1 + 0.2 * torch.randn(D, generator=g)).to(torch.bfloat16)
    shifted = oracle.shift_labels(inp["labels"])
    upstream = float((inp["labels"] != -100).sum())

    # oracle: fp32 autograd through the restated reference ops
    hp = h_pre.float().requires_grad_(True)
    sc = scale.float().requires_grad_(True)
    ww = inp["W"].float().requires_grad_(True)
    normed = torch.nn.functional.rms_norm(hp, (D,), weight=sc, eps=1e-6)
    ref = oracle.lmhead_ce(normed, ww, shifted)
    (ref * upstream).backward()

    dev = torch.device("cuda")
    emb = torch.nn.Embedding(V, D, dtype=torch.bfloat16, device=dev)
    norm = ssi_b200.RMSNormB200(D, dtype=torch.bfloat16, device=dev)
    with torch.no_grad():
        emb.weight.copy_(inp["W"])
        norm.scale.copy_(scale)
    loss_fn = ssi_b200.FusedLinearCrossEntropyLoss(emb, mode=mode)
    loss_fn.set_fused_norm(norm)
    loss_fn.grad_hint = torch.tensor([upstream], device=dev)
    hc = h_pre.to(dev).requires_grad_(True)
    loss = loss_fn(hc, shifted.to(dev))
    (loss * upstream).backward()
    torch.cuda.synchronize()
    assert abs(loss.item() - ref.item()) <= LOSS_RTOL * abs(ref.item())
    assert_grad_close(hc.grad, hp.grad, "d(pre-norm h)")
    assert_grad_close(emb.weight.grad, ww.grad, "dW", elem_roundings=DW_ELEM_ROUNDINGS)
    assert norm.scale.grad.dtype == torch.bfloat16
    assert_grad_close(norm.scale.grad, sc.grad, "dscale")
    with torch.inference_mode():
        l_eval = loss_fn(h_pre.to(dev), shifted.to(dev))
    assert abs(l_eval.item() - ref.item()) <= LOSS_RTOL * abs(ref.item())


def test_next_rows_against_the_reference_fixture():
    """The CUDA kernels against outputs of the REFERENCE's own code (tests/golden/next_rows/, generated in the build
    container by oracle/gen_golden.py; the GPU box has no reference checkout)."""
    import os

    from ssi_b200 import ops

    fx = torch.load(os.path.join(os.path.dirname(__file__), "golden", "next_rows", "next_rows_small.pt"))
    r = fx["rms"]
    y, rstd = ops.rmsnorm_fwd(r["x"].cuda(), r["scale"].cuda(), r["eps"])
    dx, dscale = ops.rmsnorm_bwd(r["dy"].cuda(), r["x"].cuda(), r["scale"].cuda(), rstd)
    _ulp_close(y, r["y"], "y")
    _ulp_close(dx, r["dx"], "dx", exact_frac=0.98)
    assert torch.allclose(dscale.cpu(), r["dscale"].float(), rtol=2 ** -6, atol=1e-4)
    t = fx["types"]
    labels = t["tokens"].clone()
    _, _, types = ops.label_shift_count_types(labels.cuda(), t["tokens"].cuda(), t["ranges"], t["pad"])
    assert dict(zip(list(t["ranges"]) + ["total"], types.tolist())) == t["counts"]
    s = fx["scale_grads"]
    g = s["grad"].cuda()
    sq = ops.grad_scale_sqnorm(g, s["scaler"].reshape(1).cuda())
    assert torch.equal(g.cpu(), s["scaled"])
    assert abs(sq.item() - s["sqnorm"].item()) <= 1e-5 * s["sqnorm"].item()
