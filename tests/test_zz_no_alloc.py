"""`fr_dedup_static` — the de-dup entry point that brings its own scratch — against `fr_dedup`.

Who needs it: a host that calls the C ABI without any device allocator to lend scratch from, a serving process
whose memory budget is already owned by another pool, or a loop where cudaMalloc/cudaFree are not allowed
because they stall every stream. The test drives both entry points through the C ABI (ctypes on the shared
library, raw pointers, raw stream handle), twice in a row with different inputs, because the built-in scratch
keeps whatever the previous call left in it. Outputs are integers, so equality is exact.

This file sorts last on purpose: a run that stops at the first failure has already run everything else.
"""
import ctypes

import pytest
import torch

from fedrec_b200.ops import native

pytestmark = pytest.mark.gpu


def _lib():
    lib = ctypes.CDLL(native.native_lib_path())
    vp, ci = ctypes.c_void_p, ctypes.c_int
    lib.fr_dedup.restype = ci
    lib.fr_dedup.argtypes = [vp, ci, ci, vp, vp, vp, vp, vp]
    lib.fr_dedup_static.restype = ci
    lib.fr_dedup_static.argtypes = [vp, ci, ci, vp, vp, vp, vp]
    lib.fr_dedup_static_max_rows.restype = ci
    lib.fr_dedup_workspace_ints.restype = ctypes.c_longlong
    lib.fr_dedup_workspace_ints.argtypes = [ci]
    lib.fr_last_error.restype = ctypes.c_char_p
    return lib


def _run(lib, fn_static, ids, N, dev):
    K = ids.numel()
    uniq = torch.full((K,), -1, dtype=torch.int32, device=dev)
    inv = torch.full((K,), -1, dtype=torch.int32, device=dev)
    count = torch.full((1,), -1, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    if fn_static:
        rc = lib.fr_dedup_static(ids.data_ptr(), K, N, uniq.data_ptr(), count.data_ptr(), inv.data_ptr(), stream)
    else:
        ws = torch.empty(lib.fr_dedup_workspace_ints(N), dtype=torch.int32, device=dev)
        rc = lib.fr_dedup(ids.data_ptr(), K, N, ws.data_ptr(), uniq.data_ptr(), count.data_ptr(), inv.data_ptr(), stream)
    assert rc == 0, lib.fr_last_error().decode()
    torch.cuda.synchronize(dev)
    return uniq, count, inv


def test_static_scratch_dedup_matches_caller_scratch_dedup():
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    lib = _lib()
    assert lib.fr_dedup_static_max_rows() == 1 << 20
    g = torch.Generator(device=dev).manual_seed(2024)
    # two calls in a row; the second uses a smaller table and other ids, so stale scratch would show
    for K, N, hi in [(2750, 65536, 9000), (4100, 30000, 700)]:
        ids = torch.randint(0, hi, (K,), device=dev, generator=g, dtype=torch.int32)
        u0, c0, i0 = _run(lib, False, ids, N, dev)
        u1, c1, i1 = _run(lib, True, ids, N, dev)
        assert int(c1) == int(c0) == torch.unique(ids).numel()
        assert torch.equal(u1, u0) and torch.equal(i1, i0)
        assert torch.equal(u1.long()[i1.long()], ids.long())


def test_static_scratch_dedup_refuses_oversized_tables():
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    lib = _lib()
    ids = torch.zeros(8, dtype=torch.int32, device=dev)
    out = torch.zeros(8, dtype=torch.int32, device=dev)
    cnt = torch.zeros(1, dtype=torch.int32, device=dev)
    rc = lib.fr_dedup_static(ids.data_ptr(), 8, (1 << 20) + 1, out.data_ptr(), cnt.data_ptr(), out.data_ptr(),
                             torch.cuda.current_stream(dev).cuda_stream)
    assert rc != 0
    msg = lib.fr_last_error().decode()
    assert "exceed" in msg and "1048576" in msg
