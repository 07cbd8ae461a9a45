"""CPU: the tensor-core lowering (stride-phase TMA views, sub-pixel phases of
the transposed conv, OOB padding) is pure integer logic in tc_plan.cu; the
debug hook executes that plan on the host and must reproduce torch."""
import ctypes as C

import pytest
import torch
import torch.nn.functional as F

from abi import make_desc
from b200conv3d._lib import ConvDesc

CASES = [
    (False, 16, 16, 4, 2, 1, 1, (8, 8, 8)), (False, 16, 16, 3, 1, 1, 1, (5, 6, 7)),
    (False, 16, 16, 3, 1, 2, 2, (7, 7, 7)), (False, 16, 16, 3, 1, 4, 4, (9, 9, 9)),
    (False, 16, 16, 1, 2, 0, 1, (6, 6, 6)), (False, 24, 16, 3, 2, 0, 1, (9, 9, 9)),
    (False, 16, 16, 3, 2, 1, 1, (7, 8, 9)), (False, 16, 24, 4, 1, 0, 1, (4, 4, 4)),
    (True, 16, 16, 4, 2, 1, 1, (4, 4, 4)), (True, 16, 16, 4, 1, 0, 1, (1, 1, 1)),
    (True, 16, 16, 3, 2, 0, 1, (5, 5, 5)), (True, 16, 16, 3, 1, 1, 1, (5, 4, 3)),
    (True, 16, 16, 1, 2, 0, 1, (3, 3, 3)),
]


@pytest.fixture(scope="module")
def dbg(product_path):
    lib = C.CDLL(product_path)
    lib.b2c_debug_plan_run.restype = C.c_int
    lib.b2c_debug_plan_run.argtypes = [C.POINTER(ConvDesc), C.c_int] + [C.c_void_p] * 4
    return lib


@pytest.mark.parametrize("case", CASES)
def test_plan_matches_torch(dbg, case):
    tr, cin, cout, k, s, p, d, sp = case
    torch.manual_seed(2)
    x = torch.randn(2, cin, *sp, requires_grad=True)
    w = torch.randn(cin, cout, k, k, k) if tr else torch.randn(cout, cin, k, k, k)
    y = F.conv_transpose3d(x, w, None, s, p, 0, 1, d) if tr else F.conv3d(x, w, None, s, p, d)
    g = torch.randn_like(y)
    gx, = torch.autograd.grad(y, x, g)
    desc = make_desc(tr, tuple(x.shape), tuple(y.shape), (k,) * 3, (s,) * 3, (p,) * 3, (d,) * 3, torch.bfloat16)
    T = k ** 3
    nd = lambda t: t.detach().permute(0, 2, 3, 4, 1).contiguous()
    if tr:   # torch [cin][cout][T]
        wf, wd = w.reshape(cin, cout, T).permute(2, 1, 0).contiguous(), w.reshape(cin, cout, T).permute(2, 0, 1).contiguous()
    else:    # torch [cout][cin][T]
        wf, wd = w.reshape(cout, cin, T).permute(2, 0, 1).contiguous(), w.reshape(cout, cin, T).permute(2, 1, 0).contiguous()
    for pas, inp, wp, ref in ((0, nd(x), wf, nd(y)), (1, nd(g), wd, nd(gx))):
        out = torch.full(ref.shape, float("nan"))
        info = (C.c_int32 * 8)()
        rc = dbg.b2c_debug_plan_run(C.byref(desc), pas, inp.data_ptr(), wp.data_ptr(), out.data_ptr(), info)
        assert rc == 0
        assert info[3] * info[4] * info[5] * info[6] in (128, 256, 512)   # 1, 2 or 4 UMMA M sub-tiles
        assert torch.isfinite(out).all()                              # every output voxel written
        assert (out - ref).abs().max() < 2e-4 * ref.abs().max()


def test_plan_rejects_unaligned_channels(dbg):
    desc = make_desc(False, (1, 3, 4, 4, 4), (1, 16, 4, 4, 4), (3,) * 3, (1,) * 3, (1,) * 3, (1,) * 3, torch.bfloat16)
    assert dbg.b2c_debug_plan_run(C.byref(desc), 0, None, None, None, None) == -2
