"""GPU: the allocation-free entry point b2c_channel_stats_static (static __device__ scratch, no
counterpart in the reference) against the existing b2c_channel_stats, through the C ABI.

Collected last on purpose (file name)."""
import ctypes as C

import pytest
import torch

from b200conv3d import _lib

pytestmark = pytest.mark.gpu

STATIC_LIMIT = 4 << 20            # B2C_STATIC_SCRATCH_BYTES (include/b2c.h)
ERR_STATIC_LIMIT = -6


def _p(t):
    return C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _stats_ws(lib, x):
    rows, c = x.shape
    dt = 0 if x.dtype == torch.bfloat16 else 2
    ws = torch.empty(max(int(lib.b2c_stats_workspace(rows, c)), 16), dtype=torch.uint8, device="cuda")
    s, q = torch.empty(c, device="cuda"), torch.empty(c, device="cuda")
    rc = lib.b2c_channel_stats(_p(x), rows, c, dt, _p(s), _p(q), _p(ws), ws.numel(), _stream())
    assert rc == 0, lib.b2c_strerror(rc)
    return s, q


def _stats_static(lib, x):
    rows, c = x.shape
    dt = 0 if x.dtype == torch.bfloat16 else 2
    # outputs are allocated by the caller BEFORE the call; the call itself allocates nothing
    s, q = torch.full((c,), float("nan"), device="cuda"), torch.full((c,), float("nan"), device="cuda")
    rc = lib.b2c_channel_stats_static(_p(x), rows, c, dt, _p(s), _p(q), _stream())
    assert rc == 0, lib.b2c_strerror(rc)
    return s, q


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_static_scratch_matches_workspace_entry_point_twice_in_a_row(product_path, dtype):
    lib = _lib._bind(product_path)
    torch.manual_seed(7)
    # two calls in a row with DIFFERENT shapes: the second must not see the first call's partials
    cases = [(32 * 16 ** 3, 128), (4 * 8 ** 3 + 3, 40)]
    xs = [(torch.randn(r, c, device="cuda") * 2 + 0.5).to(dtype) for r, c in cases]
    got = [_stats_static(lib, x) for x in xs]            # back to back on one stream
    ref = [_stats_ws(lib, x) for x in xs]
    torch.cuda.synchronize()
    for (s, q), (rs, rq), x in zip(got, ref, xs):
        # same kernels, same partition, same order: the tolerance of the BatchNorm tests (1e-4 of max) holds
        # with room to spare; in fact the results are bit-identical
        assert torch.equal(s, rs) and torch.equal(q, rq)
        xf = x.double()
        assert ((s.double() - xf.sum(0)).abs().max() / xf.sum(0).abs().max()).item() < 1e-4
        assert ((q.double() - (xf * xf).sum(0)).abs().max() / (xf * xf).sum(0).abs().max()).item() < 1e-4
    # and once more on the first shape (the array now holds the second call's contents)
    s, q = _stats_static(lib, xs[0])
    torch.cuda.synchronize()
    assert torch.equal(s, ref[0][0]) and torch.equal(q, ref[0][1])


def test_static_scratch_allocates_nothing_and_rejects_oversize(product_path):
    lib = _lib._bind(product_path)
    x = torch.randn(4096, 64, device="cuda").to(torch.bfloat16)
    _stats_static(lib, x)                               # per-device set-up happens on the first call
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    s, q = torch.empty(64, device="cuda"), torch.empty(64, device="cuda")
    free1, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        assert lib.b2c_channel_stats_static(_p(x), 4096, 64, 0, _p(s), _p(q), _stream()) == 0
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free1 <= free0   # no device allocation by the library
    # beyond the static maximum: a clear error, nothing launched
    rows, c = 1 << 20, 2048
    assert lib.b2c_stats_workspace(rows, c) > STATIC_LIMIT
    n0 = lib.b2c_launch_count()
    rc = lib.b2c_channel_stats_static(_p(x), rows, c, 0, _p(s), _p(q), _stream())
    assert rc == ERR_STATIC_LIMIT and b"static scratch" in lib.b2c_strerror(rc)
    assert lib.b2c_launch_count() == n0
