"""LaggedReader: values come back in order, `depth` pushes late, nothing is lost."""
import torch

import layerwise_b200  # noqa: F401
from layerwise_b200.utils import LaggedReader


def test_values_arrive_in_order_one_step_late():
    r = LaggedReader(depth=1)
    got = []
    for i in range(5):
        v = r.push(torch.tensor(float(i)))
        if v is not None:
            got.append(v)
    assert got == [0.0, 1.0, 2.0, 3.0]
    assert r.drain() == [4.0]
    assert r.drain() == []


def test_depth_zero_is_an_immediate_read_and_slots_are_reused():
    r = LaggedReader(depth=0)
    assert [r.push(torch.tensor(2.5)), r.push(torch.tensor(3.5))] == [2.5, 3.5]
    assert len(r._free) == 1
