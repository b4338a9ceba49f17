"""Host-side logic that needs no GPU: the packed gradient block's layout and the span a multi-GPU step
all-reduces (splatloc_b200/_C.py), used by bench.py and splatloc_b200/graphs.py."""
import pytest
import torch

from splatloc_b200 import _C


@pytest.mark.parametrize("P,C,M", [(1, 4, 0), (5, 3, 16), (1001, 4, 0), (4096, 3, 4), (500000, 4, 0)])
def test_trainable_span_is_the_leading_blocks_of_the_gradient_allocation(P, C, M):
    widths = [3, 1, 3, 4, C, 3 * M, 3, 3, 1, 4, 6]
    blocks = _C._grad_block(P, widths, "cpu")
    flat = blocks[0]._base
    assert all(b._base is flat or b.numel() == 0 for b in blocks)
    assert all(b.storage_offset() % 4 == 0 for b in blocks)            # 16-byte aligned blocks
    assert [b.numel() for b in blocks] == [P * w for w in widths]
    n = _C.trainable_grad_numel(P, C, M)
    assert n == blocks[6].storage_offset()                               # everything before dL/dmean2D
    span = blocks[0].as_strided((n,), (1,), blocks[0].storage_offset())
    span.fill_(1.0)
    assert all(bool((b == 1).all()) for b in blocks[:6]) and all(bool((b == 0).all()) for b in blocks[6:])
    assert float(flat.sum()) == float(n)
