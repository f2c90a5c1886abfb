"""HBM-roofline measurement of the rows SURVEY.md §8(f) marks "next" plus the HBM-bound kernels of the path itself
(embedding scatter-add, forward-only loss), at the sizes of one configuration.  CUDA events, L2 flushed between
iterations by the size of the buffers themselves (all > 126 MB except where noted).

    python tools/bench_rows.py [--config c3] [--iters 20]      -> one JSON object on stdout
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "speech-integration_b200"))

import torch  # noqa: E402

from ssi_b200 import _lib, ops, synth  # noqa: E402
from ssi_b200.step import PathStep  # noqa: E402


def timed(fn, iters, flush):
    for _ in range(3):
        fn()
    e0 = [torch.cuda.Event(enable_timing=True) for _ in range(iters)]
    e1 = [torch.cuda.Event(enable_timing=True) for _ in range(iters)]
    for i in range(iters):
        flush.zero_()  # 512 MB write: evicts the 126 MB L2
        e0[i].record()
        fn()
        e1[i].record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in zip(e0, e1))
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c3")
    ap.add_argument("--iters", type=int, default=20)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    _lib.load()
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(
        os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0}
    hbm = peaks["hbm_gbs"]
    cfg = synth.CONFIGS[args.config]
    B, S, V, D = cfg.batch, cfg.seq, cfg.vocab, synth.HIDDEN
    N = B * S
    g = torch.Generator().manual_seed(synth.SEED)
    tokens_h, labels_h = synth.make_tokens_labels(cfg, g)
    tokens, labels = tokens_h.to(dev), labels_h.to(dev)
    gd = torch.Generator(device=dev).manual_seed(1)
    X = torch.randn(N, D, generator=gd, device=dev).to(torch.bfloat16)
    dY = (torch.randn(N, D, generator=gd, device=dev) * 1e-2).to(torch.bfloat16)
    W = (torch.randn(V, D, generator=gd, device=dev) * 0.02).to(torch.bfloat16)
    scale = torch.ones(D, dtype=torch.bfloat16, device=dev)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    rows = {}

    def add(name, ms, nbytes, note=""):
        rows[name] = {"ms": round(ms, 4), "algorithmic_bytes": int(nbytes), "GBps": round(nbytes / ms / 1e6, 1),
                      "frac_of_hbm_peak": round(nbytes / ms / 1e6 / hbm, 3), "note": note}

    y, rstd = ops.rmsnorm_fwd(X, scale)
    add("rmsnorm_fwd", timed(lambda: ops.rmsnorm_fwd(X, scale), args.iters, flush), N * D * 4 + N * 4,
        "read x, write y (bf16) + rstd")
    add("rmsnorm_bwd", timed(lambda: ops.rmsnorm_bwd(dY, X, scale, rstd), args.iters, flush), N * D * 6 + N * 4,
        "read dy, x, write dx (bf16); dscale partials stay in L2")
    ranges = {"text": (0, 127999), "special": (128000, 128255), "dsu": (128256, V - 3), "modality": (V - 2, V - 1)}
    add("label_shift_count_types", timed(lambda: ops.label_shift_count_types(labels, tokens, ranges, 128004),
                                         args.iters, flush), N * 24, "read labels + tokens, write labels (int64); "
        "launch-latency bound at this size")
    grad = torch.randn(V, D, generator=gd, device=dev).to(torch.bfloat16)
    sc = torch.full((1,), 1.0, device=dev)
    add("grad_scale_sqnorm", timed(lambda: ops.grad_scale_sqnorm(grad, sc), args.iters, flush), V * D * 4,
        "read + write the tied gradient once (reference: 2 passes, scale then norm)")
    add("embedding_gather", timed(lambda: ops.embedding_gather(W, tokens), args.iters, flush), N * D * 4 + N * 8,
        "read rows of W, write [N, D]")
    dWg = torch.zeros(V, D, dtype=torch.bfloat16, device=dev)
    uniq = int(torch.unique(tokens).numel())
    add("embedding_grad_scatter", timed(lambda: ops.embedding_grad_scatter(dY, tokens.reshape(-1), dWg), args.iters,
                                        flush), N * D * 2 + N * 8 + uniq * D * 4,
        f"K3: read dH + ids, RMW of the {uniq} touched rows (sort included)")
    ids = tokens.reshape(-1)
    add("embedding_grad_sort", timed(lambda: ops.embedding_grad_sort(ids, V, D), args.iters, flush), N * 8 + N * 12,
        "K3 sort alone (one block): read ids, write sorted keys + rows + run index")
    sort_ws = ops.embedding_grad_sort(ids, V, D)
    add("embedding_grad_scatter_sorted", timed(lambda: ops.embedding_grad_scatter_sorted(dY, sort_ws, dWg),
                                               args.iters, flush), N * D * 2 + N * 12 + uniq * D * 4,
        "K3 segmented reduce alone (order given): read dH rows, RMW of the touched rows")
    path = PathStep(W)
    Xb = X.reshape(B, S, D)
    ms = timed(lambda: path.forward_only(Xb, labels), args.iters, flush)
    n_valid = int((labels_h[:, 1:] != synth.IGNORE).sum())
    out = {"config": args.config, "rows": N, "vocab": V, "hidden": D, "hbm_peak_GBps": hbm, "kernels": rows,
           "eval_forward_only": {"ms": round(ms, 3), "tokens_per_s": round(N / ms * 1e3),
                                 "algorithmic_tflops": round(2.0 * n_valid * V * D / ms / 1e9, 1),
                                 "note": "ssi/eval.py:27-33 path: label shift + K1 online-LSE only, nothing but lse [N] "
                                         "written"}}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
