"""Host -> device feeding for the path step: what ``batch_to_device`` + a pinned-memory DataLoader do for the
reference trainer (ssi/trainer.py:386-387), with one batch of lookahead.

``DeviceFeeder`` copies batch i+1 from pinned host memory on a side stream while batch i is being processed, into
one of two reusable sets of device buffers; ``LossReader`` brings every step's scalar result back through a pinned
buffer without stalling the launch of the next step (it reads step i-1 while step i runs).  PyTorch streams and
events only: plumbing, no arithmetic.
"""
from __future__ import annotations

from typing import Dict, Iterable, Iterator

import torch


class DeviceFeeder:
    def __init__(self, batches: Iterable[Dict[str, torch.Tensor]], device: torch.device, depth: int = 2):
        self.it = iter(batches)
        self.device = device
        self.depth = depth
        self.copy_stream = torch.cuda.Stream(device)
        self.slots = [dict() for _ in range(depth)]      # name -> device buffer
        self.ready = [torch.cuda.Event() for _ in range(depth)]
        self.released = [None] * depth                   # event: compute finished with the slot
        self.bytes_per_batch = 0

    def _issue(self, slot: int) -> bool:
        try:
            host = next(self.it)
        except StopIteration:
            return False
        bufs = self.slots[slot]
        n = 0
        with torch.cuda.stream(self.copy_stream):
            if self.released[slot] is not None:
                self.copy_stream.wait_event(self.released[slot])
            for k, h in host.items():
                if not h.is_pinned():
                    raise RuntimeError(f"ssi_b200.DeviceFeeder: host tensor {k!r} must be in pinned memory")
                d = bufs.get(k)
                if d is None or d.shape != h.shape or d.dtype != h.dtype:
                    d = torch.empty(h.shape, dtype=h.dtype, device=self.device)
                    bufs[k] = d
                d.copy_(h, non_blocking=True)
                n += h.numel() * h.element_size()
            self.ready[slot].record(self.copy_stream)
        self.bytes_per_batch = n
        return True

    def __iter__(self) -> Iterator[Dict[str, torch.Tensor]]:
        cur = torch.cuda.current_stream(self.device)
        pending = []
        for s in range(self.depth - 1):
            if self._issue(s):
                pending.append(s)
        nxt = (self.depth - 1) % self.depth
        while pending:
            slot = pending.pop(0)
            if self._issue(nxt):  # lookahead copy overlaps the compute of `slot`
                pending.append(nxt)
            nxt = (nxt + 1) % self.depth
            cur.wait_event(self.ready[slot])
            yield self.slots[slot]
            ev = torch.cuda.Event()
            ev.record(cur)
            self.released[slot] = ev


class LossReader:
    """push(loss) enqueues an async D2H copy of a 0-dim device scalar; it returns the PREVIOUS step's value (a
    Python float, or None on the first call); flush() returns the last one."""

    def __init__(self, device: torch.device):
        self.device = device
        self.host = [torch.empty(1, dtype=torch.float32).pin_memory() for _ in range(2)]
        self.ev = [None, None]
        self.i = 0

    def _get(self, j):
        if self.ev[j] is None:
            return None
        self.ev[j].synchronize()
        self.ev[j] = None
        return float(self.host[j][0])

    def push(self, loss: torch.Tensor):
        j = self.i & 1
        prev = self._get(j ^ 1)
        self.host[j].copy_(loss.detach().reshape(1), non_blocking=True)
        e = torch.cuda.Event()
        e.record(torch.cuda.current_stream(self.device))
        self.ev[j] = e
        self.i += 1
        return prev

    def flush(self):
        return self._get((self.i - 1) & 1) if self.i else None
