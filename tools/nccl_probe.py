"""Tuning helper: time NCCL all-reduce of the tied gradient buffer (one process per GPU under torchrun)."""
import os, torch, torch.distributed as dist
r = int(os.environ["LOCAL_RANK"]); torch.cuda.set_device(r)
dist.init_process_group("nccl", device_id=torch.device("cuda", r))
V, D = 133258, 2048
for dt in (torch.bfloat16, torch.float32):
    x = torch.ones(V, D, dtype=dt, device="cuda")
    for nb in (1, 4, 16):
        rows = (V + nb - 1) // nb
        for _ in range(3):
            for r0 in range(0, V, rows): dist.all_reduce(x[r0:r0 + rows])
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            for r0 in range(0, V, rows): dist.all_reduce(x[r0:r0 + rows])
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        if dist.get_rank() == 0:
            print(f"{dt} buckets={nb}: {ms:.3f} ms  algbw {x.numel()*x.element_size()/ms/1e6:.1f} GB/s", flush=True)
        x.fill_(1)
c = torch.ones(2, dtype=torch.int64, device="cuda")
torch.cuda.synchronize(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
for _ in range(3): dist.all_reduce(c)
e0.record()
for _ in range(20): dist.all_reduce(c)
e1.record(); torch.cuda.synchronize()
if dist.get_rank() == 0: print(f"tiny all_reduce: {e0.elapsed_time(e1)/20*1e3:.1f} us", flush=True)
dist.destroy_process_group()
