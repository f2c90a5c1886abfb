"""Time the sparse (embedding-part) exchange of the tied gradient alone, on the bench's own c5 token ids: ranks are
aligned by the exchange itself; per-class device times (scatter = compact segmented reduce, exchange = dp_sparse_add)
from the library's launch timers, distinct ids per rank and in the union, bytes pulled over NVLink.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/sparse_probe.py [--iters 20]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "speech-integration_b200"))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from ssi_b200 import _lib, ops, synth  # noqa: E402
from ssi_b200 import dist as ddist  # noqa: E402
from ssi_b200.peer import TiedGradExchange  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--config", default="c5")
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    _lib.load()
    cfg = synth.CONFIGS[args.config]
    V, D = cfg.vocab, synth.HIDDEN
    b0, b1 = ddist.shard_rows(cfg.batch, rank, world)
    tokens_h, _, _, dH0 = synth.global_batch_rows(cfg, b0, b1, dev)
    ids = tokens_h.to(dev).reshape(-1)
    dH = dH0.reshape(-1, D).contiguous()
    N = ids.numel()
    ex = TiedGradExchange(V, D, dist.group.WORLD, dev, max_rows=N)
    dW = torch.zeros(V, D, dtype=torch.bfloat16, device=dev)
    sort_ws = ops.embedding_grad_sort(ids, V, D)
    uniq = torch.unique(ids)
    u_local = int(uniq.numel())
    present = torch.zeros(V, dtype=torch.int32, device=dev)
    present[uniq] = 1
    cnt = present.clone()
    dist.all_reduce(cnt)
    u_union = int((cnt > 0).sum())
    pulled = int(cnt.sum()) - u_local  # rows this rank reads from its peers
    for _ in range(3):
        ex.embedding_exchange(dH, sort_ws, dW)
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    _lib.profile_begin()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.iters):
        ex.embedding_exchange(dH, sort_ws, dW)
    e1.record()
    torch.cuda.synchronize()
    cls = _lib.profile_end()
    row = {"rank": rank, "rows": N, "distinct_ids": u_local, "ms_per_call": e0.elapsed_time(e1) / args.iters,
           "compact_reduce_ms": cls["scatter"][0] / args.iters, "sparse_add_ms": cls["exchange"][0] / args.iters,
           "rows_pulled": pulled, "MB_pulled": pulled * D * 2 / 1e6}
    rows = [None] * world
    dist.all_gather_object(rows, row)
    if rank == 0:
        print({"world": world, "ids_in_union": u_union, "histogram_of_presence": torch.bincount(cnt[cnt > 0]).tolist()})
        for r in rows:
            print(r, flush=True)
    ex.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
