"""Synthetic inputs of BASELINE.json's five configurations.  Seeds follow the reference
(``SEED = 42_831``, ssi/constants.py:12; rank offset as tt/torchtune/training/seed.py:58).

Vocabulary layout (ssi/extend_llama3_2/__init__.py:94-101): ``[0,128000)`` text, ``[128000,128000+n_dsus)`` DSUs,
2 modality tokens, 256 Llama-3 specials (BOS/EOS/pad among them)  =>  V = 128256 + n_dsus + 2
(ssi/llama_configs.py:98-99).
"""
from __future__ import annotations

from dataclasses import dataclass

import torch

SEED = 42_831
HIDDEN = 2048
N_TEXT = 128_000
IGNORE = -100


@dataclass(frozen=True)
class PathConfig:
    name: str
    n_dsus: int
    batch: int
    seq: int
    kind: str  # "hubert_cpt" | "sft" | "cpt_interleaved" | "cpt_packed"
    desc: str

    @property
    def vocab(self) -> int:
        return N_TEXT + 256 + self.n_dsus + 2

    @property
    def rows(self) -> int:
        return self.batch * self.seq


CONFIGS = {
    "c1": PathConfig("c1", 5000, 1, 2048, "hubert_cpt",
                     "Llama-3.2-1B-5000-dsus tied LM head + chunked CE, 1x2048 HuBERT DSU tokens"),
    "c2": PathConfig("c2", 1024, 8, 1024, "sft",
                     "SFT SpeechTokenizer n_dsus=1024, batch 8x1024, ignore_index prompt masking"),
    "c3": PathConfig("c3", 2048, 16, 2048, "cpt_interleaved", "CPT interleaved Mimi n_dsus=2048, batch 16x2048"),
    "c4": PathConfig("c4", 8192, 8, 8192, "cpt_packed", "CPT concatenated FocalCodec n_dsus=8192, packed 8x8192"),
    "c5": PathConfig("c5", 5000, 64, 2048, "hubert_cpt", "HuBERT 5000-dsu CPT data-parallel batch 64x2048"),
}


def _special_base(cfg: PathConfig) -> int:
    return N_TEXT + cfg.n_dsus + 2


def make_tokens_labels(cfg: PathConfig, g: torch.Generator, batch: int | None = None, seq: int | None = None,
                       vocab: int | None = None):
    """tokens, labels int64 [B, S] (CPU). ``vocab`` < real V rescales id ranges for small test problems."""
    B = batch or cfg.batch
    S = seq or cfg.seq
    V = vocab or cfg.vocab
    n_text = N_TEXT if vocab is None else max(1, int(V * 0.6))
    n_dsu = cfg.n_dsus if vocab is None else max(1, int(V * 0.3))
    spec0 = n_text + n_dsu  # modality + specials live in [spec0, V)
    n_spec = max(1, V - spec0)
    bos, eos, pad = spec0 + min(2, n_spec - 1), spec0 + min(3, n_spec - 1), spec0 + min(4, n_spec - 1)

    def text(n):
        return torch.randint(0, n_text, (n,), generator=g)

    def dsu(n):
        return n_text + torch.randint(0, n_dsu, (n,), generator=g)

    tokens = torch.empty(B, S, dtype=torch.int64)
    labels = torch.empty(B, S, dtype=torch.int64)
    for b in range(B):
        if cfg.kind in ("hubert_cpt", "cpt_interleaved"):
            row = []
            total = 1
            row.append(torch.tensor([bos]))
            is_dsu = True
            while total < S:
                if cfg.kind == "hubert_cpt":
                    ln = int(torch.randint(40, 400, (1,), generator=g)) if is_dsu else int(
                        torch.randint(10, 120, (1,), generator=g))
                else:
                    # word-span lengths ~ max(Binomial(39, 0.1), 1) words (ssi/data/cpt.py:158-160), ~4 tokens/word
                    words = max(int(torch.binomial(torch.tensor(39.0), torch.tensor(0.1), generator=g)), 1)
                    ln = words * (12 if is_dsu else 2)
                seg = dsu(ln) if is_dsu else text(ln)
                row.append(torch.tensor([spec0 + (0 if is_dsu else min(1, n_spec - 1))]))  # modality token
                row.append(seg)
                total += ln + 1
                is_dsu = not is_dsu
            t = torch.cat(row)[:S]
            tokens[b] = t
            labels[b] = t  # labels = tokens.copy() (ssi/data/cpt.py:153)
        elif cfg.kind == "sft":
            P = int(torch.randint(S // 4, 3 * S // 4 + 1, (1,), generator=g))
            L = int(torch.randint(min(P + 32, S), S + 1, (1,), generator=g))
            t = torch.full((S,), pad, dtype=torch.int64)
            t[0] = bos
            t[1:P] = text(P - 1)
            t[P:L - 1] = dsu(max(L - 1 - P, 0))
            t[L - 1] = eos
            lab = t.clone()
            lab[:P] = IGNORE  # prompt masked (ssi/data/sft.py:222-228)
            lab[L - 1] = IGNORE  # EOS masked (tt/torchtune/models/llama3/_tokenizer.py:318-319)
            lab[L:] = IGNORE  # padding (ssi/data/__init__.py:174-183)
            tokens[b], labels[b] = t, lab
        elif cfg.kind == "cpt_packed":
            t = torch.full((S,), pad, dtype=torch.int64)
            lab = torch.full((S,), IGNORE, dtype=torch.int64)
            pos = 0
            tail = int(torch.randint(0, 65, (1,), generator=g))
            while pos < S - tail:
                ln = int(torch.randint(256, 1025, (1,), generator=g))
                ln = min(ln, S - tail - pos)
                nt = ln // 3
                doc = torch.cat([torch.tensor([bos]), text(max(nt - 1, 0)), dsu(max(ln - nt - 1, 0)),
                                 torch.tensor([eos])])[:ln]
                t[pos:pos + doc.numel()] = doc
                lab[pos:pos + doc.numel()] = doc
                pos += doc.numel()
            tokens[b], labels[b] = t, lab
        else:
            raise ValueError(cfg.kind)
    return tokens, labels


def make_inputs(cfg: PathConfig, rank: int = 0, batch: int | None = None, seq: int | None = None,
                vocab: int | None = None, hidden: int = HIDDEN, seed: int = SEED):
    """dict of CPU tensors: W [V,D] bf16 ~N(0,0.02^2), X [B,S,D] bf16 ~N(0,1), dH0 [B,S,D] bf16 ~N(0,1e-3^2),
    tokens/labels int64 [B,S].  W uses the global seed (replicated across ranks); the rest uses seed + rank."""
    V = vocab or cfg.vocab
    B = batch or cfg.batch
    S = seq or cfg.seq
    gw = torch.Generator().manual_seed(seed)
    W = (torch.randn(V, hidden, generator=gw) * 0.02).to(torch.bfloat16)
    g = torch.Generator().manual_seed(seed + 1 + rank)
    tokens, labels = make_tokens_labels(cfg, g, B, S, vocab)
    X = torch.randn(B, S, hidden, generator=g).to(torch.bfloat16)
    dH0 = (torch.randn(B, S, hidden, generator=g) * 1e-3).to(torch.bfloat16)
    return {"W": W, "X": X, "dH0": dH0, "tokens": tokens, "labels": labels}


def global_batch_rows(cfg: PathConfig, b0: int, b1: int, device, hidden: int = HIDDEN, seed: int = SEED):
    """Sequences ``[b0, b1)`` of the GLOBAL batch of ``cfg`` (generated once with the global seed
    and sliced contiguously by rank): tokens / labels come from one global CPU generator and are sliced; X and dH0 are
    drawn on ``device`` with one generator seed PER SEQUENCE (seed + 1000 + b), so every world size sees the same
    global batch without any rank materialising all of it.  Returns ``(tokens, labels)`` on the CPU (int64) and
    ``(X, dH0)`` on the device (bf16)."""
    g = torch.Generator().manual_seed(seed + 1)
    tokens, labels = make_tokens_labels(cfg, g)
    gd = torch.Generator(device=device)
    X = torch.empty(b1 - b0, cfg.seq, hidden, dtype=torch.bfloat16, device=device)
    dH0 = torch.empty_like(X)
    for b in range(b0, b1):
        gd.manual_seed(seed + 1000 + b)
        X[b - b0] = torch.randn(cfg.seq, hidden, generator=gd, device=device).to(torch.bfloat16)
        dH0[b - b0] = (torch.randn(cfg.seq, hidden, generator=gd, device=device) * 1e-3).to(torch.bfloat16)
    return tokens[b0:b1].contiguous(), labels[b0:b1].contiguous(), X, dH0
