#!/usr/bin/env python
"""Benchmark of the fused tied-LM-head + chunked-CE path (BASELINE.json metric) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config c5]

One "step" = one pass of the hot path over one batch of synthetic input, driven through the drop-in
API (ssi_b200.install via PathStep) in the CAUSAL order of a training step: embedding gather (+ the id sort its
backward needs), label shift + token counts, the fused projection + CE forward and backward (default single-pass
mode: per row chunk K1 logits tiles -> bf16 block + online LSE, softmax-grad in place, dW_acc += G^T X, dX = G W,
then rescale/round in backward; --mode recompute: K1 online-LSE forward, K2 recompute backward), and only then --
where the decoder backward would have produced dH0 -- the K3 tied-embedding scatter-add.  For N > 1 the tied
gradient is exchanged in three parts (token counts through the peers' ctrl blocks, the LM-head part behind the last dX
contraction, the embedding part as compact rows after the embedding backward; --exchange nccl = bf16 all-reduce after
the embedding backward).  The default workload is BASELINE.json configs[4] / SURVEY c5 — HuBERT 5000-DSU vocabulary
(V = 133,258, the "@133k vocab" the metric is quoted on), global batch 64 x 2048, generated ONCE with the global seed
and sliced by sequences over the ranks (strong scaling: the global batch is fixed; at N = 1 the whole batch runs on
one GPU).

Prints ONE JSON line on rank 0.  `value` = global token positions per second with inputs resident in HBM;
`e2e` = the same through the public module API with pinned-host inputs copied in and the loss read back every
step; `roofline` = the dominant kernel class timed live with CUDA events on its own stream; `cpu_baseline` = the
CPU restatement of the reference path (oracle/, bf16 like the reference) on a bounded sample.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "speech-integration_b200"))

METRIC = "fused LM-head+CE fwd/bwd tokens/sec @133k vocab"
UNIT = "tokens/s"
CPU_SAMPLE_ROWS = 2048  # 1 x 2048 tokens = BASELINE.json configs[0] (the reference's own CPU-runnable case)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="c5", help="SURVEY.md §8d configuration (c1..c5)")
    ap.add_argument("--mode", default=None, choices=["single_pass", "recompute"],
                    help="gradient mode of the fused loss (default: the library default, single_pass)")
    ap.add_argument("--exchange", default=None, choices=["peer", "nvls", "nccl"],
                    help="N > 1: how the tied gradient is summed over ranks (default: peer-memory reduce behind "
                         "the last dX contraction; nccl = bf16 all_reduce after the step)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-fwd-only", action="store_true")
    ap.add_argument("--no-gpu-reference", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="N > 1: skip rank 0's N = 1 re-run of the global batch")
    ap.add_argument("--no-fold", action="store_true", help="skip the scale_grads / grad-norm fold block (SURVEY 8f-3)")
    return ap.parse_args()


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return {"bf16_tflops": p["bf16_tflops"], "bf16_tflops_sustained": p.get("bf16_tflops_sustained"),
                "hbm_gbs": p["hbm_gbs"], "source": "measured (MEASURED_PEAKS.json)"}
    return {"bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "hbm_gbs": 6650.0,
            "source": "fallback (B200_PROFILING.md)"}


# ----------------------------------------------------------------------------------------- clocks sampling
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""

    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines: list[str] = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits", "-lms", "200",
                 "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
                power.append(float(f[3]))
            except ValueError:
                continue
            for name, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(power) if power else None}


# ----------------------------------------------------------------------------------------- CPU reference arm
def cpu_reference_run(cfg, iters: int, warmup: int, dtype_name: str = "bf16"):
    """The reference's CPU path (oracle restatement; bf16 is the reference's dtype, conf/common.yaml:32) on a bounded
    sample of the workload: 1 x 2048 token rows at the workload's vocabulary, fwd + bwd + embedding backward.
    Returns (tokens/s from the MEDIAN iteration, times, cores, last outputs)."""
    import statistics

    import torch

    from oracle import reference_path as oracle
    from ssi_b200 import synth

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    dt_ = torch.bfloat16 if dtype_name == "bf16" else torch.float32
    inp = synth.make_inputs(cfg, batch=1, seq=CPU_SAMPLE_ROWS)
    times, out = [], None
    for i in range(warmup + iters):
        t0 = time.perf_counter()
        out = oracle.train_step_path(inp["tokens"], inp["labels"], inp["X"], inp["dH0"], inp["W"], compute_dtype=dt_)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    return CPU_SAMPLE_ROWS / statistics.median(times), times, cores, (inp, out)


def cpu_sample_text(cfg, dtype_name, times, warmup):
    return (f"{CPU_SAMPLE_ROWS} token rows (1x2048) of {cfg.name} at V={cfg.vocab}, D=2048, {dtype_name}, fwd+bwd+"
            f"embedding backward per step; median of {len(times)} iterations after {warmup} warm-up")


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import statistics

    import torch

    from ssi_b200 import synth

    cfg = synth.CONFIGS[args.config]
    # each step = one bounded sample; same warm-up count as our arm, steps capped so the run ends within minutes
    steps = max(1, min(args.steps, 20))
    warm = max(args.warmup, 3)
    tps, times, cores, _ = cpu_reference_run(cfg, steps, warm, "bf16")
    ms = 1e3 * statistics.median(times)
    sample = cpu_sample_text(cfg, "bf16", times, warm)
    line = {
        "impl": "reference", "metric": METRIC, "value": tps, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": {"workload": workload_name(cfg)},
        "reference_arm": "CPU restatement of the reference path (oracle/reference_path.py: same torch ops at the "
                         "reference's call sites), host cores only",
        "parallel_info": torch.__config__.parallel_info().strip().splitlines()[:6],
        "cpu_baseline": {"value": tps, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": tps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------- GPU contender
def gpu_reference_tokens_per_s(W, tokens, labels, X, dH0, iters: int = 3):
    """"The kernel to beat" (SURVEY.md §2.2 / §8d): the reference's own op sequence on the SAME GPU, eager PyTorch in
    bf16 -- F.embedding, F.linear per chunk (tied_linear.py:25; cuBLAS), upcast + F.cross_entropy(sum) per chunk
    (ce_chunked_output_loss.py:41-43; ATen), (loss * n).backward(), embedding backward.  A contender, not product
    code: restated here with torch ops only, nothing of ssi_b200 runs in it."""
    import statistics

    import torch
    import torch.nn.functional as F

    Wp = W.detach().clone().requires_grad_(True)
    n_chunks, ignore = 8, -100
    shifted = torch.hstack((labels[..., 1:], torch.full_like(labels[..., -1:], ignore)))
    n_iter = (labels != ignore).sum()
    times = []
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for i in range(iters + 1):
        Wp.grad = None
        e0.record()
        h0 = F.embedding(tokens, Wp)
        h = X.detach().requires_grad_(True)
        logits = [F.linear(c, Wp) for c in h.chunk(n_chunks, dim=1)]
        total = (shifted != ignore).sum()
        loss = 0.0
        for lg, lb in zip(logits, shifted.chunk(n_chunks, dim=1)):
            loss = loss + F.cross_entropy(lg.reshape(-1, lg.size(-1)).float(), lb.reshape(-1), ignore_index=ignore,
                                          reduction="sum")
        del logits
        ((loss / total) * n_iter).backward()
        h0.backward(dH0)
        e1.record()
        torch.cuda.synchronize()
        if i > 0:
            times.append(e0.elapsed_time(e1))
    del Wp
    torch.cuda.empty_cache()
    rows = tokens.numel()
    return rows / (statistics.median(times) * 1e-3), times


def exchange_is_foldable(path) -> bool:
    return path.sync is None or path.sync.peer is not None


def workload_name(cfg) -> str:
    return (f"{cfg.name}: {cfg.desc}; V={cfg.vocab}, D=2048, global batch {cfg.batch}x{cfg.seq} "
            f"= {cfg.rows} token rows, bf16 in / fp32 accumulate")


# ----------------------------------------------------------------------------------------- our arm
def grad_checksums(g):
    """Two order-sensitive integer checksums of a bf16 tensor's bit pattern (exact, device-side)."""
    import torch

    v = g.view(torch.int16).reshape(-1).to(torch.int64)
    w = (torch.arange(v.numel(), device=v.device, dtype=torch.int64) % 8191) + 1
    return torch.stack([v.sum(), (v * w).sum()])


def oracle_parity(path, out, tokens, labels, X, dH0, W, world, dev, ignore):
    """Parity of the TIMED shape against the oracle at any world size: the closed-form fp32 check of
    tests/test_gpu_full_size.py, evaluated on the GPU tensors with the oracle's own functions where they exist
    (oracle.tied_linear: bf16 logits like tied_linear.py:25; oracle.shift_labels) --
      loss    global sum of the per-row CE on bf16 logits / global valid count * num_tokens_iter vs the summed
              loss_batch of the ranks (ssi/trainer.py:391-395; tt/recipes/full_finetune_distributed.py:844-851,867);
      dX      256 sampled token rows of THIS rank: (softmax - onehot) * scale rounded to bf16, times W, in fp32;
      dW      256 vocabulary rows (half of them label ids): every rank forms its fp32 partial G^T X over its rows plus
              its embedding rows (sum of dH0 rows per token id) with plain torch; the [256, D] partials are
              all-reduced and compared with weight.grad -- the tied gradient the exchange produced.
    Errors are reported as the worst |err| / (2^-7 |ref| + 2^-8 max|ref|) (<= 1 passes assert_grad_close) and the
    relative Frobenius error.  Every rank runs it (collectives inside); rank 0 reports."""
    import torch
    import torch.distributed as dist

    from oracle import reference_path as oracle

    V, D = W.shape
    Xf = X.reshape(-1, D)
    n_local = Xf.size(0)
    shifted = oracle.shift_labels(labels).reshape(-1)
    valid = shifted != ignore
    cnt = torch.stack([(labels != ignore).sum(), valid.sum()]).to(torch.float64)
    if world > 1:
        dist.all_reduce(cnt)
    n_iter, n_shift = cnt[0].item(), cnt[1].item()
    scale = n_iter / max(n_shift, 1.0)
    # ---- one chunked pass of the oracle's forward: per-row logsumexp and the CE sum on bf16 logits
    lse = torch.zeros(n_local, dtype=torch.float32, device=dev)
    ce_sum = torch.zeros(1, dtype=torch.float64, device=dev)
    with torch.no_grad():
        for r0 in range(0, n_local, 4096):
            sl = slice(r0, min(r0 + 4096, n_local))
            logits = oracle.tied_linear(Xf[sl], W).float()  # ce_chunked_output_loss.py:41 upcasts the bf16 logits
            l = torch.logsumexp(logits, dim=-1)
            lse[sl] = l
            v = valid[sl]
            z = logits.gather(1, shifted[sl].clamp_min(0)[:, None])[:, 0]
            ce_sum += ((l - z) * v).sum().to(torch.float64)
            del logits
        lb = out[0].detach().reshape(1).to(torch.float64).clone()
        if world > 1:
            dist.all_reduce(ce_sum)
            dist.all_reduce(lb)
        want_loss = ce_sum.item() / max(n_shift, 1.0) * n_iter
        g = torch.Generator().manual_seed(7)

        def worst(got, ref):
            got, ref = got.float(), ref.float()
            bound = ref.abs() * 2.0 ** -7 + ref.abs().max() * 2.0 ** -8
            return float(((got - ref).abs() / bound).max().item()), rel_fro(got, ref)

        # ---- dX: sampled token rows of this rank
        rows = torch.randperm(n_local, generator=g)[:256].to(dev)
        lg = oracle.tied_linear(Xf[rows], W).float()
        G = torch.softmax(lg, dim=-1)
        rv = valid[rows]
        G[torch.arange(rows.numel(), device=dev)[rv], shifted[rows][rv]] -= 1.0
        G = (G * scale * rv[:, None]).to(torch.bfloat16).float()  # autograd hands bf16 dlogits to the matmuls
        dx_over, dx_fro = worst(out[1].reshape(-1, D)[rows], G @ W.float())
        dxs = torch.tensor([dx_over, dx_fro], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dxs, op=dist.ReduceOp.MAX)
        # ---- dW: the same vocabulary rows on every rank (label ids of rank 0 + random ids)
        labs = shifted[valid][:128]
        labs = torch.cat([labs, torch.zeros(128 - labs.numel(), dtype=labs.dtype, device=dev)])
        if world > 1:
            dist.broadcast(labs, src=0)
        vs = torch.unique(torch.cat([labs, torch.randint(0, V, (128,), generator=g).to(dev)]))
        Xv = Xf[valid].float()
        Gv = torch.exp((Xv @ W[vs].float().t()).to(torch.bfloat16).float() - lse[valid][:, None])
        Gv -= (shifted[valid][:, None] == vs[None, :]).float()
        part = (Gv * scale).to(torch.bfloat16).float().t() @ Xv
        del Xv, Gv
        hit = (tokens.reshape(-1)[:, None] == vs[None, :]).float()
        part += hit.t() @ dH0.reshape(-1, D).float()  # backward of nn.Embedding: the rows of dH0 summed per id
        del hit
        if world > 1:
            dist.all_reduce(part)
        dw_over, dw_fro = worst(path.weight.grad[vs], part)
    return {"loss_times_n": float(lb.item()), "loss_times_n_oracle": want_loss,
            "loss_rel_err": abs(float(lb.item()) - want_loss) / abs(want_loss),
            "dX_worst_over_bound": float(dxs[0].item()), "dX_rel_frobenius": float(dxs[1].item()),
            "dW_worst_over_bound": dw_over, "dW_rel_frobenius": dw_fro, "rows": 256 * world,
            "vocab_rows": int(vs.numel()), "rows_checked_shape": f"{n_local} token rows per rank x {world} ranks",
            "how": "closed-form fp32 oracle on the timed shape: loss over ALL rows, 256 sampled dX rows per rank (max "
                   "over ranks), tied dW (LM head + embedding) on sampled vocabulary rows from all-reduced fp32 "
                   "partials; bound = 2^-7 |ref| + 2^-8 max|ref| (tests/test_gpu_parity.py), <= 1 passes"}


def rel_fro(a, b):
    return float(((a.float() - b.float()).norm() / (b.float().norm() + 1e-30)).item())


def run_ours(args):
    import torch
    import torch.distributed as dist

    from ssi_b200 import _lib, synth
    from ssi_b200 import dist as ddist
    from ssi_b200.step import PathStep

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the ssi_b200 path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    pg = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        pg = dist.group.WORLD
    _lib.load()
    peaks = load_peaks()
    cfg = synth.CONFIGS[args.config]
    D = synth.HIDDEN
    V = cfg.vocab
    warmup = max(args.warmup, 3)

    # ---- synthetic inputs: W replicated (global seed); the GLOBAL batch is generated once with the global seed and
    # this rank takes its contiguous slice of sequences, so losses are comparable across N
    b0, b1 = ddist.shard_rows(cfg.batch, rank, world)
    Bl = b1 - b0
    gw = torch.Generator().manual_seed(synth.SEED)
    W = (torch.randn(V, D, generator=gw) * 0.02).to(torch.bfloat16).to(dev)
    tokens_h, labels_h, X, dH0 = synth.global_batch_rows(cfg, b0, b1, dev)
    tokens, labels = tokens_h.to(dev), labels_h.to(dev)
    n_valid_local = int((labels_h[:, 1:] != synth.IGNORE).sum())
    nv = torch.tensor([n_valid_local], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(nv)
    n_valid_global = int(nv.item())

    path = PathStep(W, process_group=pg, mode=args.mode, exchange=args.exchange, max_tokens_per_rank=Bl * cfg.seq)
    mode = path.loss_fn.mode

    def step():
        path.zero_grad(set_to_none=True)  # ssi/trainer.py:410
        return path(tokens, labels, X, dH0)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        step()
    barrier()
    # short steps (N = 8: 21 ms) leave 3 warm-up steps far too brief for clocks, allocator and event pools to settle:
    # keep warming up, untimed, until one second of steps has run (the count actually used is what the line reports)
    t_w = time.perf_counter()
    step()
    torch.cuda.synchronize()
    est = torch.tensor([time.perf_counter() - t_w], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(est, op=dist.ReduceOp.MAX)  # every rank runs the same number of extra steps
    extra = max(0, min(60, int(1.0 / max(est.item(), 1e-3)) - warmup))
    for _ in range(extra):
        step()
    warmup_run = warmup + 1 + extra
    barrier()

    # ---- timed region: EXACTLY K steps, CUDA events, max over ranks
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = _lib.launch_count()
    torch.cuda.reset_peak_memory_stats(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    marks = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    e0.record()
    t_host0 = time.perf_counter()
    for i in range(args.steps):
        out = step()
        marks[i].record()
    host_ms_step = (time.perf_counter() - t_host0) * 1e3 / args.steps  # time to ENQUEUE a step (no sync inside)
    e1.record()
    barrier()
    step_ms = [round((e0 if i == 0 else marks[i - 1]).elapsed_time(marks[i]), 3) for i in range(args.steps)]
    ms_total = e0.elapsed_time(e1)
    launches = _lib.launch_count() - launches0
    peak_alloc = torch.cuda.max_memory_allocated(dev)
    # ---- the same K steps once more with the library's per-launch profiler on (two event records around each of the
    # ~90 launches of a step): the kernel classes and the roofline come from THIS pass, the headline from the
    # un-instrumented one above; both ms_per_step are reported so that the cost of the events is a number
    _lib.profile_begin()
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    p0.record()
    for i in range(args.steps):
        out = step()
    p1.record()
    barrier()
    classes = _lib.profile_end()
    clocks = sampler.stop() if rank == 0 else None
    tp = torch.tensor([p0.elapsed_time(p1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tp, op=dist.ReduceOp.MAX)
    ms_step_prof = tp.item() / args.steps
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = t.item() / args.steps
    value = cfg.rows / (ms_step * 1e-3)
    loss_batch = out[0].detach().clone()

    # ---- parity block: the global loss * n and the tied gradient against the N = 1 path on the SAME global batch
    # (N > 1; rank 0 re-runs the whole batch alone), the gradient's bit pattern identical on every rank
    parity = None
    if world > 1:
        lg = loss_batch.reshape(1).to(torch.float64)
        dist.all_reduce(lg)
        cs = grad_checksums(path.weight.grad)
        all_cs = [torch.empty_like(cs) for _ in range(world)]
        dist.all_gather(all_cs, cs)
        same = all(bool(torch.equal(c, all_cs[0])) for c in all_cs)
        parity = {"loss_times_n_global": float(lg.item()), "grad_bits_identical_on_all_ranks": same}
        if rank == 0 and not args.no_parity:
            tk, lb, Xg, dHg = synth.global_batch_rows(cfg, 0, cfg.batch, dev)
            solo = PathStep(W.detach().clone())
            l1, _, _ = solo(tk.to(dev), lb.to(dev), Xg, dHg)
            torch.cuda.synchronize()
            parity.update({
                "loss_times_n_at_n1": float(l1.item()),
                "loss_rel_err_vs_n1": abs(float(lg.item()) - float(l1.item())) / abs(float(l1.item())),
                "dW_rel_frobenius_vs_n1": rel_fro(path.weight.grad, solo.weight.grad),
                "how": "rank 0 re-ran the whole 64 x 2048 global batch alone (N = 1 path, same W) after the timed "
                       "region; N = 1 itself is checked against the CPU oracle in its own bench line and in tests/"})
            del solo, Xg, dHg
            from ssi_b200 import ops as _ops

            _ops.free_workspaces()
            torch.cuda.empty_cache()
        barrier()

    # ---- oracle parity on the timed shape, at every world size (all ranks take part)
    oracle_blk = None
    if not args.no_parity:
        try:
            oracle_blk = oracle_parity(path, out, tokens, labels, X, dH0, W, world, dev, synth.IGNORE)
        except Exception as e:  # noqa: BLE001 - report, do not lose the bench line
            oracle_blk = {"error": f"{type(e).__name__}: {str(e)[:200]}"}
        torch.cuda.empty_cache()
        barrier()

    # ---- forward only (the eval path, ssi/eval.py:27-33): K1 alone
    fwd_only = None
    if not args.no_fwd_only:
        for _ in range(2):
            path.forward_only(X, labels)
        barrier()
        kf = max(3, min(args.steps, 10))
        e0.record()
        for _ in range(kf):
            lf = path.forward_only(X, labels)
        e1.record()
        barrier()
        tf = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tf, op=dist.ReduceOp.MAX)
        ms_f = tf.item() / kf
        fwd_only = {"value": cfg.rows / (ms_f * 1e-3), "unit": UNIT, "ms_per_step": ms_f, "steps": kf,
                    "algorithmic_tflops_per_gpu": 2.0 * n_valid_local * V * D / (ms_f * 1e-3) / 1e12,
                    "loss_rank0": float(lf.item())}

    # ---- scale_grads + the tied weight's grad norm folded into the step, against the step followed by
    # the library's stand-alone one-pass kernel (the reference runs two passes, ssi/trainer.py:404-408)
    fold_blk = None
    if not args.no_fold and mode == "single_pass" and exchange_is_foldable(path):
        try:
            from ssi_b200 import ops as _ops

            kf = max(3, min(args.steps, 5))
            # the scaler of ssi/trainer.py:404 with one micro-batch per step: 1 / num_tokens_iter (UNSHIFTED labels,
            # :391), summed over ranks
            nit = torch.tensor([int((labels_h != synth.IGNORE).sum())], dtype=torch.int64, device=dev)
            if world > 1:
                dist.all_reduce(nit)
            scaler = torch.full((1,), 1.0 / max(int(nit.item()), 1), device=dev)

            def timed(fn):
                fn()
                barrier()
                e0.record()
                for _ in range(kf):
                    fn()
                e1.record()
                barrier()
                tt = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
                if world > 1:
                    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                return tt.item() / kf

            def two_pass():
                step()
                return _ops.grad_scale_sqnorm(path.weight.grad, scaler)

            ms_two = timed(two_pass)
            sq_two = float(two_pass().item())
            path.set_fold(True)

            def folded():
                step()
                return path.fold.sqnorm()

            ms_fold = timed(folded)
            sq_fold = float(folded().item())
            path.set_fold(False)
            fold_blk = {"ms_per_step_folded": ms_fold, "ms_per_step_plus_standalone_pass": ms_two,
                        "saved_ms_per_step": ms_two - ms_fold, "grad_sqnorm_folded": sq_fold,
                        "grad_sqnorm_standalone": sq_two, "sqnorm_rel_diff": abs(sq_fold - sq_two) / max(sq_two, 1e-30),
                        "steps": kf, "what": "step + scale_grads(1/num_tokens) + sum(g^2) of the tied weight: folded "
                        "into apply_dw / peer reduce + embedding backward vs one extra pass over the [V, D] gradient"}
        except Exception as e:  # an optional block must never cost the bench line
            fold_blk = {"error": f"{type(e).__name__}: {str(e)[:200]}"}
            try:
                path.set_fold(False)
            except Exception:
                pass

    # ---- end to end through the public API with pinned host inputs
    e2e = None
    if not args.no_e2e:
        hx, hd = X.cpu().pin_memory(), dH0.cpu().pin_memory()
        ht, hl = tokens_h.pin_memory(), labels_h.pin_memory()
        h2d = sum(x.numel() * x.element_size() for x in (hx, hd, ht, hl))

        from ssi_b200.feed import DeviceFeeder, LossReader

        host_batch = {"tokens": ht, "labels": hl, "hidden": hx, "dH0": hd}

        def e2e_run(n):
            """n steps through the public API: pinned host batch -> (one-batch lookahead copy on a side stream)
            -> PathStep -> loss read back on the host every step (one step behind the launches)."""
            reader = LossReader(dev)
            losses = []
            for b in DeviceFeeder((host_batch for _ in range(n)), dev):
                path.zero_grad(set_to_none=True)
                lb, _, _ = path(b["tokens"], b["labels"], b["hidden"], b["dH0"])
                prev = reader.push(lb)
                if prev is not None:
                    losses.append(prev)
            losses.append(reader.flush())
            return losses

        e2e_run(2)
        barrier()
        k2 = max(3, min(args.steps, 10))
        e0.record()
        e2e_losses = e2e_run(k2)
        e1.record()
        barrier()
        assert len(e2e_losses) == k2
        t2 = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t2, op=dist.ReduceOp.MAX)
        e2e = {"value": cfg.rows / (t2.item() / k2 * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": 4, "steps": k2, "ms_per_step": t2.item() / k2,
               "how": "pinned host batch -> DeviceFeeder (H2D of step i+1 overlaps step i) -> PathStep -> loss D2H "
                      "every step (LossReader, read one step behind)", "loss_last": e2e_losses[-1]}
        del hx, hd

    # per-rank sum of the main-stream kernel classes: under a power cap the GPUs of one box differ by a few per
    # cent, and every step ends at the slowest rank (the exchange is a barrier)
    per_rank_ms = None
    if world > 1:
        mine = sum(ms for k, (ms, _) in classes.items() if k not in ("exchange",)) / args.steps
        per_rank_ms = [None] * world
        dist.all_gather_object(per_rank_ms, round(mine, 3))
    if path.sync is not None:
        path.sync.check(confirm_multiplier=True)  # raises if the exchange ran with a multiplier that backward did not confirm / timed out
    exchange = path.exchange
    if path.sync is not None and path.sync.peer is not None and path.sync.peer.multicast_ptr:
        exchange = "peer+nvls (multimem.ld_reduce in the dense reduce)"
    path.close()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel class (timed live above with events on the launching stream)
    alg_total = 6.0 * n_valid_local * V * D  # this rank: fwd 2NVD + dX 2NVD + dW 2NVD on valid rows only
    alg = {"lse_fwd": 2.0 * n_valid_local * V * D, "dlogits": 0.0, "dW": 2.0 * n_valid_local * V * D,
           "dX": 2.0 * n_valid_local * V * D}
    n_local = Bl * cfg.seq
    # executed flop: compacted rows are padded to 128-row tiles; recompute mode runs the logits contraction twice
    executed = {k: 2.0 * n_valid_local * V * D for k in ("lse_fwd", "dlogits", "dW", "dX")}
    kern = {}
    for k, (ms, cnt) in classes.items():
        per_step = ms / args.steps
        kern[k] = {"ms_per_step": per_step, "launches_per_step": cnt / args.steps,
                   "share_of_step": per_step / ms_step_prof}
        if k in executed and per_step > 0:
            kern[k]["executed_tflops"] = executed[k] / (per_step * 1e-3) / 1e12
            kern[k]["algorithmic_tflops"] = alg[k] / (per_step * 1e-3) / 1e12
    dom = max((k for k in ("lse_fwd", "dW", "dX")), key=lambda k: classes[k][0])
    dom_ms = classes[dom][0] / max(classes[dom][1], 1)
    dom_alg = alg[dom] * args.steps / max(classes[dom][1], 1)
    achieved = dom_alg / (dom_ms * 1e-3) / 1e12
    peak = peaks["bf16_tflops_sustained"] or peaks["bf16_tflops"]
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get(dom)
        except Exception:
            traffic = None
    roofline = {"bound": "tensor", "kernel": f"gemm_kernel<{dom}>", "achieved": achieved, "peak": peak,
                "unit": "TFLOP/s", "frac": achieved / peak, "traffic": traffic,
                "traffic_source": "static: one ncu --set full capture of this kernel at this shape "
                                  "(profiles/traffic.json), bytes per launch; not re-measured in this run",
                "peak_kind": "sustained bf16 cuBLAS (kernel timed inside a long step), " + peaks["source"],
                "frac_of_burst_peak": achieved / peaks["bf16_tflops"], "burst_peak": peaks["bf16_tflops"],
                "algorithmic_flops_per_launch": dom_alg, "avg_launch_ms": dom_ms}
    # the HBM-bound row the north star singles out: K3 embedding scatter (sort + segmented reduce, one record per call)
    k3 = None
    if classes["scatter"][1] > 0 and world == 1:
        uniq = int(torch.unique(tokens_h).numel())
        k3_bytes = n_local * D * 2 + n_local * 8 + 2 * uniq * D * 2
        k3 = {"bound": "hbm", "algorithmic_bytes_per_step": k3_bytes,
              "ms_per_step": classes["scatter"][0] / args.steps,
              "achieved_gbs": k3_bytes / (classes["scatter"][0] / args.steps * 1e-3) / 1e9,
              "peak_gbs": peaks["hbm_gbs"],
              "frac": k3_bytes / (classes["scatter"][0] / args.steps * 1e-3) / 1e9 / peaks["hbm_gbs"],
              "unique_ids": uniq, "note": "sort (embedding forward) + segmented reduce (embedding backward)"}
    step_tflops = alg_total / (ms_step * 1e-3) / 1e12
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": warmup, "warmup_steps_run": warmup_run, "ms_per_step": ms_step,
        "ms_per_step_with_kernel_profiler": ms_step_prof, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "memory": {"peak_allocated_gib": peak_alloc / 2 ** 30,
                   "what": "torch.cuda.max_memory_allocated over the timed steps on rank 0: inputs, W, .grad, the "
                           "single-pass workspace (E block + fp32 dW_acc) and outputs; a peer exchange buffer "
                           "(cudaMalloc / symmetric memory) is outside torch's allocator and not counted"},
        "config": {"workload": workload_name(cfg), "rows_per_rank": n_local, "valid_rows_global": n_valid_global,
                   "parallelism": f"dp{world}", "mode": mode, "exchange": exchange,
                   "flow": "causal: the embedding gradient is handed over only after dX was returned "
                           "(ssi/trainer.py:393-394), through install()",
                   "l2": "inputs (X, W, dlogits blocks) are several times larger than "
                   "the 126 MB L2, no explicit flush", "loss_batch_rank0": float(loss_batch.item())},
        "clocks": clocks, "gpu_launches": launches, "host_enqueue_ms_per_step": host_ms_step,
        "step_ms_rank0": step_ms, "e2e": e2e,
        "roofline": roofline,
        "step_algorithmic_tflops_per_gpu": step_tflops,
        "step_frac_of_bf16_peak": {"burst": step_tflops / peaks["bf16_tflops"], "sustained": step_tflops / peak},
        "kernels": kern, "fwd_only": fwd_only, "parity": parity, "k3_roofline": k3, "scale_grads_fold": fold_blk,
    }
    if per_rank_ms is not None:
        line["per_rank_kernel_ms_per_step"] = per_rank_ms
    if world == 1 and not args.no_gpu_reference:
        # the reference's own op sequence (cuBLAS + ATen, eager bf16) on this GPU, on a bounded slice of the batch
        nb = min(Bl, 16)
        try:
            tps_ref, tms = gpu_reference_tokens_per_s(W, tokens[:nb], labels[:nb], X[:nb], dH0[:nb])
            line["gpu_reference"] = {
                "value": tps_ref, "unit": UNIT, "ms_per_step": sorted(tms)[len(tms) // 2],
                "rows": nb * cfg.seq, "ratio_ours_over_it": value / tps_ref,
                "what": "the reference's op sequence on this GPU, eager PyTorch bf16: F.embedding, F.linear per chunk "
                        "(cuBLAS), upcast + F.cross_entropy(sum) per chunk (ATen), backward, embedding backward; "
                        f"{nb} of the {Bl} sequences (its fp32 log-prob tensors need N*V*8 B); median of "
                        f"{len(tms)} after 1 warm-up"}
        except Exception as e:  # out of memory on a smaller part: report, do not fail the bench
            line["gpu_reference"] = {"unavailable": f"{type(e).__name__}: {str(e)[:160]}"}
            torch.cuda.empty_cache()
    if world == 1 and not args.no_cpu_baseline:
        tps, times, cores, (inp, ref) = cpu_reference_run(cfg, 3, 1, "bf16")
        line["cpu_baseline"] = {"value": tps, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": cpu_sample_text(cfg, "bf16", times, 1),
                                "parallel_info": torch.__config__.parallel_info().strip().splitlines()[:6]}
        tps32, times32, _, (inp, ref32) = cpu_reference_run(cfg, 3, 1, "fp32")
        line["cpu_baseline"]["fp32"] = {"value": tps32, "unit": UNIT, "sample": cpu_sample_text(cfg, "fp32", times32, 1)}
        # N = 1 parity: the same sample through the GPU path against the fp32 CPU oracle run just timed
        solo = PathStep(inp["W"].to(dev))
        lb, dXs, _ = solo(*(inp[k].to(dev) for k in ("tokens", "labels", "X", "dH0")))
        torch.cuda.synchronize()
        ref_loss, n_iter, ref_dH, ref_dW = ref32
        want = ref_loss.item() * n_iter
        line["parity"] = {"sample": f"{CPU_SAMPLE_ROWS} token rows of {cfg.name} (the cpu_baseline sample)",
                          "loss_times_n": float(lb.item()), "loss_times_n_oracle_fp32": want,
                          "loss_rel_err": abs(float(lb.item()) - want) / abs(want),
                          "dX_rel_frobenius": rel_fro(dXs.reshape(ref_dH.shape).cpu(), ref_dH),
                          "dW_rel_frobenius": rel_fro(solo.weight.grad.cpu(), ref_dW),
                          "tolerance": "loss rel <= 1e-3, gradients rel-Frobenius <= 4e-3 (tests/test_gpu_parity.py)"}
    if oracle_blk is not None:
        line["parity"] = dict(line.get("parity") or {}, oracle=oracle_blk)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
