"""DeviceFeeder / LossReader (host <-> device plumbing of the end-to-end path): contents and ordering."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_feeder_yields_every_batch_in_order():
    from ssi_b200.feed import DeviceFeeder, LossReader

    dev = torch.device("cuda", 0)
    host = [{"a": torch.full((1 << 20,), float(i)).pin_memory(), "b": torch.arange(8) + i} for i in range(7)]
    for h in host:
        h["b"] = h["b"].pin_memory()
    reader = LossReader(dev)
    seen, losses = [], []
    for i, b in enumerate(DeviceFeeder(iter(host), dev)):
        assert b["a"].device.type == "cuda"
        # keep the device busy so the lookahead copy really overlaps
        x = b["a"].clone()
        for _ in range(20):
            x = x * 1.0000001
        seen.append((float(b["a"][0]), int(b["b"][0])))
        prev = reader.push(b["a"].sum() / b["a"].numel())
        if prev is not None:
            losses.append(prev)
    losses.append(reader.flush())
    assert seen == [(float(i), i) for i in range(7)]
    assert losses == [float(i) for i in range(7)]


def test_feeder_rejects_pageable_memory():
    from ssi_b200.feed import DeviceFeeder

    with pytest.raises(RuntimeError):
        list(DeviceFeeder(iter([{"a": torch.zeros(4)}]), torch.device("cuda", 0)))
