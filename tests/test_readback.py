import pytest
import torch

from fedrec_b200.utils.readback import LaggedScalars


@pytest.mark.parametrize("lag", [0, 1, 3])
def test_every_value_is_delivered_in_order_on_cpu(lag):
    r = LaggedScalars("cpu", lag=lag)
    got = []
    for i in range(10):
        v = r.push(torch.tensor(float(i)))
        if v is not None:
            got.append(v)
    got += r.drain()
    assert got == [float(i) for i in range(10)]
    assert r.drain() == []
