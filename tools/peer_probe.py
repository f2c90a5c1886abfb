"""Time the peer-memory exchange of the tied gradient alone (no compute in front, ranks aligned by the exchange
itself), and NCCL's bf16 all_reduce of the same [V, D] buffer next to it.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/peer_probe.py [--iters 20]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "speech-integration_b200"))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from ssi_b200 import _lib  # noqa: E402
from ssi_b200.peer import TiedGradExchange  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--vocab", type=int, default=133258)
    ap.add_argument("--hidden", type=int, default=2048)
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=dev)
    _lib.load()
    V, D = args.vocab, args.hidden
    ex = TiedGradExchange(V, D, dist.group.WORLD, dev)
    ex.dw_acc.fill_(float(rank + 1))
    dW = torch.empty(V, D, dtype=torch.bfloat16, device=dev)
    one = torch.ones(1, dtype=torch.float32, device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def timed(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(args.iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / args.iters], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    t_peer = timed(lambda: ex.exchange(dW, one, one))
    want = float(sum(range(1, world + 1)))
    ok = bool((dW.float() == want).all().item())
    # the same exchange with the in-switch reduction (multimem.ld_reduce on the multicast mapping)
    t_nvls, ok_nvls = None, None
    try:
        exn = TiedGradExchange(V, D, dist.group.WORLD, dev, nvls=True)
    except RuntimeError as e:
        exn, why = None, str(e)
    if exn is not None:
        exn.dw_acc.fill_(float(rank + 1))
        dW.zero_()
        t_nvls = timed(lambda: exn.exchange(dW, one, one))
        ok_nvls = bool((dW.float() == want).all().item())
        exn.close()
    g = torch.ones(V, D, dtype=torch.bfloat16, device=dev)
    t_nccl = timed(lambda: dist.all_reduce(g))
    if rank == 0:
        gb = V * D * 2 / 1e9
        if t_nvls is not None:
            print(f"world {world}: NVLS reduce + bf16 gather {t_nvls:.3f} ms (result ok={ok_nvls})", flush=True)
        else:
            print(f"world {world}: NVLS unavailable: {why[:200]}", flush=True)
        print(f"world {world}: peer exchange {t_peer:.3f} ms (fp32 reduce + bf16 gather, result ok={ok}), "
              f"nccl bf16 all_reduce {t_nccl:.3f} ms, buffer {gb:.3f} GB bf16", flush=True)
    ex.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
