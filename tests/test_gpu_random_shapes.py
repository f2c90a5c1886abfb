"""Seeded sweep of ragged problem shapes through single-pass mode (every schedule feature on: CTA pairs, 256x512
tiles with tail splitting, lockstep of all three contractions, several row chunks) against the fp32 oracle.
The shapes are chosen to hit the edges the fixed cases do not: one row, fewer rows than a tile, an odd number of
128-row tiles, vocabularies below one tile and one past a tile boundary, almost everything ignored, and logits
blocks so small that a micro-batch needs dozens of launch rounds (ragged vocabulary splits in every round)."""
import random

import pytest
import torch

from oracle import reference_path as oracle
from test_gpu_parity import DW_ELEM_ROUNDINGS, LOSS_RTOL, assert_grad_close

pytestmark = pytest.mark.gpu


def _shapes():
    rng = random.Random(20261104)
    out = [(1, 8, 64, 0.0, 4608), (127, 255, 64, 0.0, 4608), (129, 257, 128, 0.5, 4608), (385, 513, 64, 0.9, 1)]
    for _ in range(20):
        n = rng.choice([rng.randint(1, 130), rng.randint(131, 900), rng.randint(901, 2500)])
        v = rng.choice([rng.randint(8, 300), rng.randint(301, 2100), rng.randint(2101, 9000)])
        d = rng.choice([64, 128, 192, 256, 512])
        out.append((n, v, d, rng.choice([0.0, 0.3, 0.97]), rng.choice([4608, 1, 2])))
    return out


@pytest.mark.parametrize("N,V,D,p_ignore,cap_mib", _shapes())
def test_single_pass_random_shape(N, V, D, p_ignore, cap_mib):
    from ssi_b200 import _lib, ops

    g = torch.Generator().manual_seed(N * 131 + V * 7 + D)
    X = torch.randn(N, D, generator=g).to(torch.bfloat16)
    W = (torch.randn(V, D, generator=g) * (D ** -0.5)).to(torch.bfloat16)
    labels = torch.randint(0, V, (N,), generator=g)
    labels[torch.rand(N, generator=g) < p_ignore] = -100
    upstream = 37.0
    _lib.set_option("g_cap_mib", cap_mib)
    try:
        h = torch.tensor([upstream], device="cuda")
        loss_out, lse, n_valid, dX, dW_acc = ops.lmhead_ce_fwd_grad(X.cuda(), W.cuda(), labels.cuda(), h)
        dX, dW = ops.lmhead_ce_apply_grad(dX, dW_acc, h, h)
        with torch.no_grad():
            fwd_only = ops.lmhead_ce_fwd(X.cuda(), W.cuda(), labels.cuda())[0]
        torch.cuda.synchronize()
    finally:
        _lib.set_option("g_cap_mib", 4608)
    nv = int((labels != -100).sum())
    assert n_valid.item() == nv
    if nv == 0:
        assert torch.isnan(loss_out[1]) and (dX == 0).all() and (dW == 0).all()
        return
    ref_loss, rdX, rdW = oracle.lmhead_ce_fwd_bwd(X.reshape(1, N, D), W, labels.reshape(1, N), upstream,
                                                  num_output_chunks=1)
    assert abs(loss_out[1].item() - ref_loss.item()) <= LOSS_RTOL * max(abs(ref_loss.item()), 1e-3)
    assert abs(fwd_only[1].item() - ref_loss.item()) <= LOSS_RTOL * max(abs(ref_loss.item()), 1e-3)
    assert_grad_close(dX, rdX.reshape(N, D), "dX")
    assert_grad_close(dW, rdW, "dW", elem_roundings=DW_ELEM_ROUNDINGS)
