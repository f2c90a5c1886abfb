"""ORACLE TOOLING — generates tests/golden/*.pt by running the REFERENCE's own modules.

Run in the build container only (needs /root/reference, which does not exist on the GPU box):

    python oracle/gen_golden.py

The reference packages cannot be imported as packages here (torchtune/__init__ needs torchao and omegaconf,
ssi.constants needs sardalign), but the three files that ARE the path import only torch, so they are loaded by
file path: tt/torchtune/modules/loss/ce_chunked_output_loss.py, tt/torchtune/modules/tied_linear.py and
ssi/loss.py (with a stub ``torchtune.modules.TransformerDecoder``, which ssi/loss.py uses as a type hint only).
Each fixture stores small seeded inputs and the reference's outputs (loss, dH, dW of the tied weight including
the nn.Embedding gradient) in the reference dtype (bf16) and in fp32.
"""
from __future__ import annotations

import importlib.util
import os
import sys
import types

import torch
from torch import nn

REF = os.environ.get("SSI_REFERENCE", "/root/reference")
TT = os.path.join(REF, "third_party", "torchtune", "torchtune")
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _load(name: str, path: str):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def load_reference():
    """Returns (CEWithChunkedOutputLoss, TiedLinear, compute_loss) straight from the reference's files."""
    ce = _load("ref_ce_chunked_output_loss", os.path.join(TT, "modules", "loss", "ce_chunked_output_loss.py"))
    tl = _load("ref_tied_linear", os.path.join(TT, "modules", "tied_linear.py"))
    if "torchtune" not in sys.modules:
        stub = types.ModuleType("torchtune")
        stub_mod = types.ModuleType("torchtune.modules")
        stub_mod.TransformerDecoder = object  # type hint only (ssi/loss.py:4,7)
        stub.modules = stub_mod
        sys.modules["torchtune"] = stub
        sys.modules["torchtune.modules"] = stub_mod
    loss = _load("ref_ssi_loss", os.path.join(REF, "ssi", "loss.py"))
    return ce.CEWithChunkedOutputLoss, tl.TiedLinear, loss.compute_loss


class TinyTiedModel(nn.Module):
    """Stand-in for TransformerDecoder around the path: embedding gather (transformer.py:628), an identity
    "decoder" that adds the externally supplied hidden state (so gradients reach the embedding), and
    chunked_output through the reference's TiedLinear (transformer.py:499-502)."""

    def __init__(self, weight: torch.Tensor, hidden: torch.Tensor, TiedLinear, num_output_chunks: int):
        super().__init__()
        self.tok_embeddings = nn.Embedding(weight.size(0), weight.size(1), dtype=weight.dtype)
        with torch.no_grad():
            self.tok_embeddings.weight.copy_(weight)
        self.output = TiedLinear(self.tok_embeddings)
        self.hidden = hidden.clone().requires_grad_(True)
        self.num_output_chunks = num_output_chunks
        self.emb_out = None

    def forward(self, tokens, mask=None, encoder_input=None, encoder_mask=None, input_pos=None):
        self.emb_out = self.tok_embeddings(tokens)
        self.emb_out.retain_grad()
        return [self.output(c) for c in self.hidden.chunk(self.num_output_chunks, dim=1)]


def run_reference(tokens, labels, hidden, dH0, weight, dtype, num_output_chunks=8):
    CE, TiedLinear, compute_loss = load_reference()
    model = TinyTiedModel(weight.to(dtype), hidden.to(dtype), TiedLinear, num_output_chunks)
    loss_fn = CE(num_output_chunks=num_output_chunks)
    num_tokens_iter = int((labels != loss_fn.ignore_index).sum().item())  # ssi/trainer.py:391
    loss = compute_loss({"tokens": tokens, "labels": labels}, model, loss_fn)
    # ssi/trainer.py:393-394, plus the gradient that would flow back through the decoder into the embedding
    (loss * num_tokens_iter + (model.emb_out * dH0.to(dtype)).sum()).backward()
    return {
        "loss": loss.detach().float(),
        "num_tokens_iter": num_tokens_iter,
        "dH": model.hidden.grad.float(),
        "dW": model.tok_embeddings.weight.grad.float(),
    }


CASES = {
    # name: (cfg, B, S, V, D, chunks)
    "sft_small": ("c2", 3, 64, 515, 64, 8),
    "cpt_small": ("c3", 2, 96, 1026, 128, 8),
    "ragged_chunks": ("c1", 2, 19, 259, 64, 8),  # s=19 -> torch.chunk gives 7 chunks
    "packed_small": ("c4", 1, 300, 2050, 64, 8),
}


def _reference_function(path: str, name: str):
    """One top-level function of a reference file whose module cannot be imported here (absent packages)."""
    import ast

    tree = ast.parse(open(path).read())
    fn = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == name)
    ns = {"Tensor": torch.Tensor, "torch": torch, "nn": nn}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), path, "exec"), ns)
    return ns[name]


def next_rows_fixture():
    """Outputs of the REFERENCE's own code for the rows SURVEY.md §8(f) marks next: RMSNorm forward + autograd
    backward (tt/torchtune/modules/rms_norm.py), count_token_types (ssi/train_utils.py:150-165) and scale_grads
    (tt/torchtune/training/_grad_scaler.py:11-31) followed by the squared norm clip_grad_norm_ would take."""
    rms = _load("ref_rms_norm", os.path.join(TT, "modules", "rms_norm.py"))
    count_token_types = _reference_function(os.path.join(REF, "ssi", "train_utils.py"), "count_token_types")
    scale_grads = _reference_function(os.path.join(TT, "training", "_grad_scaler.py"), "scale_grads")
    g = torch.Generator().manual_seed(42831)
    N, D, V = 200, 256, 1000
    x = (torch.randn(N, D, generator=g) * 2.5).to(torch.bfloat16)
    scale = (1 + 0.2 * torch.randn(D, generator=g)).to(torch.bfloat16)
    dy = (torch.randn(N, D, generator=g) * 1e-2).to(torch.bfloat16)
    norm = rms.RMSNorm(D).to(torch.bfloat16)
    with torch.no_grad():
        norm.scale.copy_(scale)
    xx = x.clone().requires_grad_(True)
    y = norm(xx)
    y.backward(dy)
    tokens = torch.randint(0, V, (6, 50), generator=g)
    ranges = {"text": (0, 599), "special": (600, 639), "dsu": (640, 997), "modality": (998, 999)}
    pad = 601
    grad = (torch.randn(V, D, generator=g) * 40).to(torch.bfloat16)
    lin = nn.Embedding(V, D).to(torch.bfloat16)
    lin.weight.grad = grad.clone()
    scaler = torch.tensor(1 / 4099)
    scale_grads(lin, scaler)
    return {
        "rms": {"x": x, "scale": scale, "dy": dy, "eps": norm.eps, "y": y.detach(), "dx": xx.grad,
                "dscale": norm.scale.grad},
        "types": {"tokens": tokens, "ranges": ranges, "pad": pad, "counts": count_token_types(tokens, ranges, pad)},
        "scale_grads": {"grad": grad, "scaler": scaler, "scaled": lin.weight.grad.clone(),
                        "sqnorm": lin.weight.grad.float().pow(2).sum()},
    }


def main():
    sys.path.insert(0, os.path.join(os.path.dirname(OUT), "..", "speech-integration_b200"))
    from ssi_b200 import synth

    os.makedirs(OUT, exist_ok=True)
    for name, (cfg, B, S, V, D, chunks) in CASES.items():
        inp = synth.make_inputs(synth.CONFIGS[cfg], batch=B, seq=S, vocab=V, hidden=D)
        fx = {"inputs": inp, "chunks": chunks, "case": (cfg, B, S, V, D)}
        for tag, dt in (("bf16", torch.bfloat16), ("fp32", torch.float32)):
            out = run_reference(inp["tokens"], inp["labels"], inp["X"], inp["dH0"], inp["W"], dt, chunks)
            # keep the fixtures small: gradients stored as bf16 for the bf16 run, fp32 for the fp32 run
            if dt == torch.bfloat16:
                out["dH"], out["dW"] = out["dH"].to(dt), out["dW"].to(dt)
            fx[tag] = out
        # the all-ignored edge: 0/0 -> NaN (ce_chunked_output_loss.py:83)
        torch.save(fx, os.path.join(OUT, f"{name}.pt"))
        print(name, "loss", fx["fp32"]["loss"].item(), "n", fx["fp32"]["num_tokens_iter"],
              os.path.getsize(os.path.join(OUT, f"{name}.pt")) // 1024, "KiB")
    os.makedirs(os.path.join(OUT, "next_rows"), exist_ok=True)
    torch.save(next_rows_fixture(), os.path.join(OUT, "next_rows", "next_rows_small.pt"))
    print("next_rows_small", os.path.getsize(os.path.join(OUT, "next_rows", "next_rows_small.pt")) // 1024, "KiB")


if __name__ == "__main__":
    main()
