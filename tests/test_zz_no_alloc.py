"""GPU test of sl_knn_static, the distCUDA2 entry point that allocates nothing.

It is meant for callers that cannot obtain scratch memory themselves (a real-time hook where an
allocation's device-wide synchronisation is unaffordable, a plugin inside a process whose device
memory belongs to another allocator).  Its scratch is one array inside the library image, shared by
all calls, so the things worth checking are: it gives exactly what sl_knn gives, it still does on a
second call (nothing stale is left behind in the shared array), and it refuses inputs larger than
the array was sized for.  Everything goes through the C ABI with raw device pointers.

Kept in a file pytest collects last: a run that stops at the first failure has then already run
every other test.  It must also pass on its own in a fresh process (first call = address look-up).
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _dist_both(lib, pts):
    P = pts.shape[0]
    stream = torch.cuda.current_stream().cuda_stream
    ref = torch.zeros(P, device="cuda")
    nbytes = int(lib.sl_knn_scratch_bytes(P))
    scratch = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    assert lib.sl_knn(P, pts.data_ptr(), ref.data_ptr(), scratch.data_ptr(), nbytes, stream) == 0
    out = torch.full((P,), -1.0, device="cuda")
    rc = lib.sl_knn_static(P, pts.data_ptr(), out.data_ptr(), stream)
    assert rc == 0, lib.sl_last_error()
    torch.cuda.synchronize()
    return out, ref


def test_static_scratch_knn_matches_sl_knn_twice_in_a_row():
    from splatloc_b200 import _capi, scenes
    lib = _capi.load()
    torch.cuda.set_device(0)
    # two calls in a row with different sizes and point sets: the second must not see the first's leftovers
    for P, seed in ((50000, 3), (4097, 4), (50000, 3)):
        pts = scenes.make_knn_points(P, seed=seed).cuda().contiguous()
        out, ref = _dist_both(lib, pts)
        # same kernels, same arithmetic, only the scratch differs: bit-identical (the tolerance the
        # other distCUDA2 tests use)
        assert torch.equal(out.view(torch.int32), ref.view(torch.int32)), \
            f"P={P}: {int((out.view(torch.int32) != ref.view(torch.int32)).sum())} of {P} values differ"
        assert float(out.min()) >= 0.0


def test_static_scratch_knn_rejects_oversized_input():
    from splatloc_b200 import _capi
    lib = _capi.load()
    P = 65536 + 1
    pts = torch.rand((P, 3), device="cuda")
    out = torch.zeros(P, device="cuda")
    rc = lib.sl_knn_static(P, pts.data_ptr(), out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == -3
    msg = lib.sl_last_error().decode()
    assert "65536" in msg and "sl_knn" in msg
    # the limit itself is accepted
    P = 65536
    pts = torch.rand((P, 3), device="cuda")
    out, ref = _dist_both(lib, pts)
    assert torch.equal(out.view(torch.int32), ref.view(torch.int32))
