"""Host-side logic that needs no GPU: module surface (same attributes / errors as the reference's loss and
TiedLinear), synthetic generators, row sharding, and the loud failure of the product path without CUDA."""
import pytest
import torch
from torch import nn

import ssi_b200
from ssi_b200 import dist as ddist
from ssi_b200 import ops, synth


def test_loss_module_surface_matches_reference():
    emb = nn.Embedding(100, 64, dtype=torch.bfloat16)
    loss = ssi_b200.FusedLinearCrossEntropyLoss(emb)
    # defaults of CEWithChunkedOutputLoss (tt/torchtune/modules/loss/ce_chunked_output_loss.py:30-33)
    assert loss.num_output_chunks == 8 and loss.ignore_index == -100
    assert loss.tied_module is emb
    # the tied weight must not be registered under the loss (state-dict keys unchanged)
    assert list(loss.state_dict().keys()) == [] and list(loss.parameters()) == []
    with pytest.raises(AttributeError, match="does not have attribute 'weight'"):
        ssi_b200.FusedLinearCrossEntropyLoss(nn.ReLU())  # same error as tied_linear.py:50-53
    with pytest.raises(AttributeError):
        ssi_b200.HiddenPassthroughOutput(nn.ReLU())


def test_no_cpu_fallback():
    emb = nn.Embedding(100, 64, dtype=torch.bfloat16)
    loss = ssi_b200.FusedLinearCrossEntropyLoss(emb)
    h = torch.randn(2, 8, 64, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        loss(h, torch.zeros(2, 8, dtype=torch.int64))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.embedding_grad_scatter(torch.zeros(4, 64, dtype=torch.bfloat16), torch.zeros(4, dtype=torch.int64),
                                   torch.zeros(10, 64, dtype=torch.bfloat16))


def test_install_swaps_output_and_embedding_without_touching_state_dict():
    class Tiny(nn.Module):
        def __init__(self):
            super().__init__()
            self.tok_embeddings = nn.Embedding(50, 16)
            self.output = lambda x: x @ self.tok_embeddings.weight.t()
            self.num_output_chunks = 0

        def set_num_output_chunks(self, n):
            self.num_output_chunks = n

    m = Tiny()
    w = m.tok_embeddings.weight
    keys = list(m.state_dict().keys())
    loss = ssi_b200.install(m)
    assert list(m.state_dict().keys()) == keys == ["tok_embeddings.weight"]
    assert m.tok_embeddings.weight is w and isinstance(m.tok_embeddings, ssi_b200.TiedEmbeddingB200)
    assert isinstance(m.output, ssi_b200.HiddenPassthroughOutput) and m.output.tied_module is m.tok_embeddings
    assert m.num_output_chunks == loss.num_output_chunks == 8  # ssi/trainer.py:303-304
    x = torch.randn(2, 3, 16)
    assert m.output(x) is x
    # CPU forward of the swapped embedding still works (plain gather)
    assert m.tok_embeddings(torch.tensor([[1, 2]])).shape == (1, 2, 16)


def test_vocab_sizes_follow_reference_formula():
    # ssi/llama_configs.py:98-99: 128000 + 256 + n_dsus + 2 modality tokens
    assert synth.CONFIGS["c1"].vocab == 133258
    assert synth.CONFIGS["c2"].vocab == 129282
    assert synth.CONFIGS["c3"].vocab == 130306
    assert synth.CONFIGS["c4"].vocab == 136450
    assert synth.CONFIGS["c5"].rows == 131072


@pytest.mark.parametrize("name", ["c1", "c2", "c3", "c4"])
def test_synth_tokens_in_range_and_deterministic(name):
    cfg = synth.CONFIGS[name]
    g1, g2 = torch.Generator().manual_seed(3), torch.Generator().manual_seed(3)
    t1, l1 = synth.make_tokens_labels(cfg, g1, batch=2)
    t2, l2 = synth.make_tokens_labels(cfg, g2, batch=2)
    assert torch.equal(t1, t2) and torch.equal(l1, l2)
    assert t1.min() >= 0 and t1.max() < cfg.vocab
    assert ((l1 == -100) | ((l1 >= 0) & (l1 < cfg.vocab))).all()
    if name == "c2":
        assert (l1 == -100).float().mean() > 0.3  # prompt + padding masked
    if name in ("c1", "c3"):
        assert (l1 != -100).all()  # labels = tokens (ssi/data/cpt.py:153)


def test_shard_rows_partitions_exactly():
    for n, w in [(64, 8), (64, 3), (5, 8), (16, 1)]:
        spans = [ddist.shard_rows(n, r, w) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(spans[i][1] == spans[i + 1][0] for i in range(w - 1))
        assert max(e - b for b, e in spans) - min(e - b for b, e in spans) <= 1


def test_install_fuse_norm_keeps_the_state_dict_and_hands_over_the_scale():
    import ssi_b200
    from ssi_b200.model_hooks import FusedNormPassthrough

    class Norm(nn.Module):  # shape of torchtune's RMSNorm (tt/torchtune/modules/rms_norm.py:23-27)
        def __init__(self, dim):
            super().__init__()
            self.normalized_shape, self.eps = (dim,), 1e-6
            self.scale = nn.Parameter(torch.ones(dim))

    class Model(nn.Module):
        def __init__(self):
            super().__init__()
            self.tok_embeddings = nn.Embedding(50, 16)
            self.norm = Norm(16)
            self.output = None

    m = Model()
    keys = set(m.state_dict())
    scale = m.norm.scale
    loss_fn = ssi_b200.install(m, fuse_norm=True)
    assert isinstance(m.norm, FusedNormPassthrough) and m.norm.scale is scale
    assert set(m.state_dict()) == keys and "norm.scale" in keys
    x = torch.randn(2, 3, 16)
    assert m.norm(x) is x
    assert loss_fn._norm_args()[0] is scale and loss_fn._norm_args()[1] == 1e-6
    assert "scale" not in dict(loss_fn.named_parameters())  # the loss does not own it


def test_grad_fold_protocol_on_the_host():
    """State machine of ssi_b200.fold.TiedGradFold and the scale_grads mirror (no kernel runs: CPU tensors)."""
    from ssi_b200 import fold as sfold

    w = nn.Parameter(torch.zeros(40, 64, dtype=torch.bfloat16))
    f = sfold.TiedGradFold(w)
    assert f.dense_args(torch.tensor([4.0]), 128) == (None, None)  # not armed: nothing folds
    f.arm()                                                         # scaler = 1 / the announced multiplier
    with pytest.raises(RuntimeError):
        f.sparse_args(128)                                          # the dense site starts a fold
    ps, ws = f.dense_args(torch.tensor([4.0]), 128)
    assert ps.item() == 0.25 and ws.dtype == torch.uint8 and ws.numel() >= 4 * 4 * 4096
    with pytest.raises(RuntimeError):
        f.sparse_args(256)                                          # more token rows than the loss saw
    ps2, ws2 = f.sparse_args(128)
    assert ps2 is ps and ws2 is ws and f.folded and not f.armed
    f.arm(torch.tensor(0.125))
    assert f.dense_args(torch.tensor([4.0]), 64)[0].item() == 0.125 and not f.folded
    f.arm()
    assert f.dense_args(torch.tensor([0.0]), 64)[0].item() == 0.0   # no valid token: scale 0, not inf

    class M(nn.Module):
        def __init__(self):
            super().__init__()
            self.a = w
            self.b = nn.Parameter(torch.ones(3))

    m, holder = M(), type("L", (), {})()
    w.grad, m.b.grad = torch.ones_like(w), torch.ones(3)
    holder.fold = f
    f.folded = True
    sfold.scale_grads(m, torch.tensor(0.5), holder)
    assert torch.equal(w.grad, torch.ones_like(w)) and torch.equal(m.b.grad, torch.full((3,), 0.5))
    f.folded = False
    sfold.scale_grads(m, torch.tensor(0.5), holder)                 # nothing folded: the plain reference behaviour
    assert torch.equal(w.grad, torch.full_like(w, 0.5)) and torch.equal(m.b.grad, torch.full((3,), 0.25))


def test_fold_workspace_size_is_device_independent():
    import ctypes

    from ssi_b200 import _lib

    b = ctypes.c_size_t(0)
    assert _lib.load().ssi_b200_grad_fold_workspace_bytes(131072, 2048, ctypes.byref(b)) == 0
    assert b.value == 4 * (2 * 4096 + 2 * (131072 // 32) * 8)
    assert _lib.load().ssi_b200_grad_fold_workspace_bytes(64, 64, ctypes.byref(b)) == 0
    assert b.value == 4 * (4 * 4096)
    assert _lib.load().ssi_b200_grad_fold_workspace_bytes(0, 64, ctypes.byref(b)) < 0


def test_install_returns_a_ce_with_chunked_output_loss_when_torchtune_is_importable(monkeypatch):
    """ssi/trainer.py:303 checks ``isinstance(self.loss_fn, CEWithChunkedOutputLoss)``: with torchtune importable the
    object install() returns must pass it.  torchtune itself cannot be imported here (torchao is missing), so the
    module path is stubbed with the REFERENCE's own class loaded by file path when the checkout is present, else with
    a stand-in of the same constructor (tt/torchtune/modules/loss/ce_chunked_output_loss.py:30-33)."""
    import importlib.util
    import os
    import sys
    import types

    ref = "/root/reference/third_party/torchtune/torchtune/modules/loss/ce_chunked_output_loss.py"
    if os.path.exists(ref):
        spec = importlib.util.spec_from_file_location("_ref_ce_chunked", ref)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        ref_cls = mod.CEWithChunkedOutputLoss
    else:
        class ref_cls(nn.Module):  # noqa: N801
            def __init__(self, num_output_chunks: int = 8, ignore_index: int = -100):
                super().__init__()
                self.num_output_chunks = num_output_chunks
                self.ignore_index = ignore_index

    tt, ttm, ttl = types.ModuleType("torchtune"), types.ModuleType("torchtune.modules"), types.ModuleType(
        "torchtune.modules.loss")
    ttl.CEWithChunkedOutputLoss = ref_cls
    tt.modules, ttm.loss = ttm, ttl
    for name, m in (("torchtune", tt), ("torchtune.modules", ttm), ("torchtune.modules.loss", ttl)):
        monkeypatch.setitem(sys.modules, name, m)

    class M(nn.Module):
        def __init__(self):
            super().__init__()
            self.tok_embeddings = nn.Embedding(50, 64, dtype=torch.bfloat16)
            self.output = None
            self.chunks = None

        def set_num_output_chunks(self, n):
            self.chunks = n

    m = M()
    loss_fn = ssi_b200.install(m, num_output_chunks=4, ignore_index=-7)
    assert isinstance(loss_fn, ref_cls) and isinstance(loss_fn, ssi_b200.FusedLinearCrossEntropyLoss)
    assert loss_fn.num_output_chunks == 4 and loss_fn.ignore_index == -7 and m.chunks == 4
    assert list(loss_fn.state_dict().keys()) == [] and loss_fn.tied_module is m.tok_embeddings


def test_global_batch_is_the_same_for_every_world_size():
    """bench.py slices ONE global batch by rank: the slices of any world size concatenate to the
    batch a single process sees, tokens and labels from the global generator, X / dH0 seeded per sequence."""
    cfg = synth.CONFIGS["c1"]
    dev = torch.device("cpu")
    tok, lab, X, dH = synth.global_batch_rows(cfg, 0, cfg.batch, dev, hidden=64)
    for world in (2, 4, 8):
        parts = []
        for r in range(world):
            b0, b1 = ddist.shard_rows(cfg.batch, r, world)
            parts.append(synth.global_batch_rows(cfg, b0, b1, dev, hidden=64))
        assert torch.equal(torch.cat([p[0] for p in parts]), tok) and torch.equal(torch.cat([p[1] for p in parts]), lab)
        assert torch.equal(torch.cat([p[2] for p in parts]), X) and torch.equal(torch.cat([p[3] for p in parts]), dH)


def test_install_rejects_option_combinations_that_have_no_kernel():
    class M(nn.Module):
        def __init__(self):
            super().__init__()
            self.tok_embeddings = nn.Embedding(50, 64, dtype=torch.bfloat16)
            self.output = None

    with pytest.raises(ValueError):
        ssi_b200.install(M(), mode="recompute", fold_scale_grads=True)      # the fold lives in single-pass kernels
    with pytest.raises(ValueError):
        ssi_b200.install(M(), replace_embedding=False, fold_scale_grads=True)
    with pytest.raises(ValueError):
        ssi_b200.install(M(), mode="bogus")


def test_uninstall_gives_the_logits_path_back():
    """VERDICT r2 (#6c): install() must be reversible.  After uninstall() the model holds the very objects it had
    before -- its own output projection (a TiedLinear in the reference, tt/torchtune/models/llama3_2/
    _component_builders.py:123), embedding module, final norm and chunk count -- so TransformerDecoder.forward returns
    logits again (tt/torchtune/modules/transformer.py:651-661); parameters were shared all along."""
    class Norm(nn.Module):
        def __init__(self):
            super().__init__()
            self.scale = nn.Parameter(torch.ones(16, dtype=torch.bfloat16))
            self.eps = 1e-6

        def forward(self, x):
            return x

    class Tiny(nn.Module):
        def __init__(self):
            super().__init__()
            self.tok_embeddings = nn.Embedding(50, 16, dtype=torch.bfloat16)
            self.norm = Norm()
            self.output = lambda x: torch.nn.functional.linear(x, self.tok_embeddings.weight)
            self.num_output_chunks = 0

        def set_num_output_chunks(self, n):
            self.num_output_chunks = n

    m = Tiny()
    emb, out, norm, w = m.tok_embeddings, m.output, m.norm, m.tok_embeddings.weight
    keys = list(m.state_dict().keys())
    ssi_b200.uninstall(m)  # never installed: nothing happens
    ssi_b200.install(m, fuse_norm=True)
    ssi_b200.install(m, replace_norm=True)  # a second install keeps the ORIGINAL parts for uninstall
    assert isinstance(m.output, ssi_b200.HiddenPassthroughOutput) and m.num_output_chunks == 8
    assert list(m.state_dict().keys()) == keys
    ssi_b200.uninstall(m)
    assert m.tok_embeddings is emb and m.output is out and m.norm is norm and m.tok_embeddings.weight is w
    assert m.num_output_chunks == 0 and list(m.state_dict().keys()) == keys
    assert not hasattr(m, "_ssi_b200_saved")
    h = torch.randn(2, 3, 16).to(torch.bfloat16)
    assert m.output(h).shape == (2, 3, 50)  # logits again, not hidden states
    ssi_b200.uninstall(m)  # idempotent
