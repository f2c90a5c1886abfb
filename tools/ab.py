"""In-process A/B of library tuning options (same box, same thermal state, interleaved rounds).

    python tools/ab.py [--config c3] [--steps 6] [--rounds 3] name=v[,name=v...] name=v[,...] ...

Each positional argument is one variant: a comma-separated list of ssi_b200_set_option pairs ("base" = defaults).
Prints per-variant mean ms/step and per-kernel-class ms (CUDA events around every launch).
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "speech-integration_b200"))

import torch  # noqa: E402

from ssi_b200 import _lib, synth  # noqa: E402
from ssi_b200.step import PathStep  # noqa: E402

DEFAULTS = {"g_cap_mib": 4608, "g_tiles": 0, "group_m_grad": 16, "group_m_dw": 1, "group_m_dx": 1, "store_hint": 0, "k1_hints": 0, "dw_hints": 0, "dx_hints": 0, "lockstep": 64,
            "cta_group": 2, "wide_dw": 2, "wide_dx": 2, "tail_split": 1, "lockstep_k1": 1, "shape_gate": 1, "exp_stash": 1, "k1_group_m": 64, "k1_store_first": 1, "sparse_two_phase": -1}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c3")
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--rounds", type=int, default=3)
    ap.add_argument("--mode", default=None)
    ap.add_argument("--batch", type=int, default=None)
    ap.add_argument("variants", nargs="+")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    _lib.load()
    for k in list(DEFAULTS):
        try:
            _lib.set_option(k, DEFAULTS[k])
        except RuntimeError:
            DEFAULTS.pop(k)
    cfg = synth.CONFIGS[args.config]
    B = args.batch or cfg.batch
    D, V = synth.HIDDEN, cfg.vocab
    gw = torch.Generator().manual_seed(synth.SEED)
    W = (torch.randn(V, D, generator=gw) * 0.02).to(torch.bfloat16).to(dev)
    g = torch.Generator().manual_seed(synth.SEED + 1)
    tokens_h, labels_h = synth.make_tokens_labels(cfg, g, batch=B)
    gd = torch.Generator(device=dev).manual_seed(synth.SEED + 1)
    X = torch.randn(B, cfg.seq, D, generator=gd, device=dev).to(torch.bfloat16)
    dH0 = (torch.randn(B, cfg.seq, D, generator=gd, device=dev) * 1e-3).to(torch.bfloat16)
    tokens, labels = tokens_h.to(dev), labels_h.to(dev)
    path = PathStep(W, mode=args.mode)

    def step():
        path.zero_grad(set_to_none=True)
        return path(tokens, labels, X, dH0)

    def apply(variant):
        for k, v in DEFAULTS.items():
            _lib.set_option(k, v)
        if variant != "base":
            for kv in variant.split(","):
                k, v = kv.split("=")
                _lib.set_option(k, int(v))

    res = {v: {"ms": [], "classes": {}} for v in args.variants}
    for v in args.variants:  # warm every variant once (first-launch attribute setup, workspace growth)
        apply(v)
        step()
    torch.cuda.synchronize()
    for _ in range(args.rounds):
        for v in args.variants:
            apply(v)
            step()
            torch.cuda.synchronize()
            _lib.profile_begin()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(args.steps):
                step()
            e1.record()
            torch.cuda.synchronize()
            cls = _lib.profile_end()
            res[v]["ms"].append(e0.elapsed_time(e1) / args.steps)
            for k, (ms, cnt) in cls.items():
                res[v]["classes"].setdefault(k, []).append(ms / args.steps)
    rows = B * cfg.seq
    for v in args.variants:
        ms = res[v]["ms"]
        mean = sum(ms) / len(ms)
        cl = " ".join(f"{k} {sum(x) / len(x):.2f}" for k, x in res[v]["classes"].items() if sum(x) > 0)
        print(f"{v:40s} ms/step {mean:7.2f} (min {min(ms):.2f} max {max(ms):.2f}) tok/s {rows / mean * 1e3:9.0f} | {cl}",
              flush=True)


if __name__ == "__main__":
    main()
