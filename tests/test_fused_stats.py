"""Statistics in the GEMM epilogue (b2c_fusion.stat_kind, include/b2c.h): the per-channel (BatchNorm3d) / per-sample
(LayerNorm) sums a convolution delivers for the normalisation layer that follows it in the reference's layer tables
(dcgan.py:47-65,161-179; utils.py:62-75; FID_ResNet.py:70-76) must equal the sums of the tensor it stored.

GPU: through the C ABI for every kernel family that accumulates them (tcgen05 direct and swapped epilogues, CTA
pairs, sub-pixel transposed conv, the point GEMM of dcgan.py:47, the one-channel gather kernel of dcgan.py:161) and
for the ones that fall back to one pass over y (fp32 / CUDA-core), on ragged extents too; and at the module level
against the same network with the fusion switched off.  CPU: the host plumbing through the oracle library."""
import ctypes as C
import os
import subprocess
import sys

import pytest
import torch
import torch.nn as nn

from b200conv3d import _lib
from b200conv3d._lib import Fusion

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BF16 = torch.bfloat16

# transposed, cin, cout, k, stride, pad, spatial, batch, dtype, kind          what it exercises
CASES = [
    (False, 64, 256, 4, 2, 1, (16, 16, 16), 4, BF16, 1),     # direct epilogue, N tile 256 (CTA pair)
    (False, 64, 256, 4, 2, 1, (16, 16, 16), 4, BF16, 2),     # per-sample, several samples per CTA
    (False, 128, 320, 3, 1, 1, (6, 7, 9), 3, BF16, 1),       # ragged extents: clipped rows, two N tiles
    (False, 128, 320, 3, 1, 1, (6, 7, 9), 3, BF16, 2),
    (False, 256, 512, 4, 2, 1, (8, 8, 8), 8, BF16, 2),       # 4^3 outputs: two samples per 128-row tile
    (False, 64, 64, 3, 1, 1, (12, 12, 12), 2, BF16, 1),      # swapped epilogue (Cout <= 128): channel = TMEM lane
    (False, 64, 96, 3, 1, 1, (5, 6, 7), 2, BF16, 1),         # swapped, ragged
    (False, 64, 64, 3, 1, 1, (12, 12, 12), 2, BF16, 2),      # swapped + per-sample: one-pass fall-back
    (True, 128, 64, 4, 2, 1, (8, 8, 8), 3, BF16, 1),         # dcgan.py:51-65 ConvT -> BatchNorm3d: 8 sub-pixel lattices
    (True, 256, 256, 4, 2, 1, (4, 4, 4), 4, BF16, 1),        # same, direct epilogue
    (True, 64, 128, 4, 1, 0, (1, 1, 1), 8, BF16, 1),         # dcgan.py:47 z -> 4^3: columns are (tap, channel)
    (False, 1, 128, 4, 2, 1, (32, 32, 32), 3, BF16, 2),      # dcgan.py:161-162 Conv(1 -> ndf) -> LayerNorm: gather kernel
    (False, 1, 64, 4, 2, 1, (10, 12, 14), 5, BF16, 2),       # same, tiles straddle samples
    (False, 40, 24, 3, 1, 1, (6, 6, 6), 2, BF16, 1),         # CUDA-core kernel: fall-back pass
    (False, 16, 32, 3, 1, 1, (6, 6, 6), 2, torch.float32, 1),  # fp32: fall-back pass
    (False, 16, 32, 3, 1, 1, (6, 6, 6), 2, torch.float32, 2),
]


def _desc(tr, cin, cout, k, s, p, sp, n, dtype):
    from abi import make_desc
    osp = [((v - 1) * s - 2 * p + k) if tr else ((v + 2 * p - k) // s + 1) for v in sp]
    return make_desc(tr, (n, cin, *sp), (n, cout, *osp), (k,) * 3, (s,) * 3, (p,) * 3, (1,) * 3, dtype), osp


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES, ids=lambda c: "%s%d-%d_k%d_%s_b%d_%s_kind%d" % ("T" if c[0] else "", c[1], c[2], c[3],
                                                                                       "x".join(map(str, c[6])), c[7],
                                                                                       "bf16" if c[8] == BF16 else "f32", c[9]))
def test_conv_epilogue_statistics_equal_the_sums_of_the_stored_output(product_path, case):
    from abi import Abi
    tr, cin, cout, k, s, p, sp, n, dtype, kind = case
    gpu = Abi(product_path, "cuda:0")
    L = gpu.lib
    torch.manual_seed(cin * 7 + cout + kind)
    desc, osp = _desc(tr, cin, cout, k, s, p, sp, n, dtype)
    x = torch.randn(n, *sp, cin, device="cuda").to(dtype)
    w = (torch.randn(cin, cout, k, k, k) if tr else torch.randn(cout, cin, k, k, k)).cuda() / (cin * k ** 3 / (8 if tr else 1)) ** 0.5
    bias = torch.randn(cout, device="cuda") * 0.3
    wp = gpu.pack(desc, 0, w)
    y = torch.empty(n, *osp, cout, dtype=dtype, device="cuda")
    y2 = torch.empty_like(y)
    nst = cout if kind == 1 else n
    st = torch.full((2, nst), float("nan"), device="cuda")
    core = (L.b2c_conv3d_workspace(C.byref(desc), 0) + 255) & ~255
    ws = torch.empty(max(core + L.b2c_conv3d_stats_workspace(C.byref(desc), kind), 16), dtype=torch.uint8, device="cuda")
    fu = Fusion(C.c_void_p(bias.data_ptr()), 2, 0.2, kind, st[0].data_ptr(), st[1].data_ptr())
    P = lambda t: C.c_void_p(t.data_ptr())
    rc = L.b2c_conv3d_fwd(C.byref(desc), P(x), P(wp), P(y), C.byref(fu), P(ws), ws.numel(), gpu._st())
    assert rc == 0, L.b2c_strerror(rc)
    fu0 = Fusion(C.c_void_p(bias.data_ptr()), 2, 0.2)
    assert L.b2c_conv3d_fwd(C.byref(desc), P(x), P(wp), P(y2), C.byref(fu0), P(ws), ws.numel(), gpu._st()) == 0
    torch.cuda.synchronize()
    assert torch.equal(y, y2)                                    # asking for statistics does not change the output
    yd = y.double()
    dims = (0, 1, 2, 3) if kind == 1 else (1, 2, 3, 4)
    ref_s, ref_q = yd.sum(dim=dims), (yd * yd).sum(dim=dims)
    # fp32 partial sums of up to a few thousand bf16 values per CTA, fp64 across CTAs
    assert ((st[0].double() - ref_s).abs().max() / ref_q.sqrt().max()).item() < 1e-4, (st[0], ref_s)
    assert ((st[1].double() - ref_q).abs().max() / ref_q.max()).item() < 1e-5, (st[1], ref_q)
    # a second call gives bit-identical statistics (deterministic tree, no atomics)
    st2 = torch.empty_like(st)
    fu2 = Fusion(C.c_void_p(bias.data_ptr()), 2, 0.2, kind, st2[0].data_ptr(), st2[1].data_ptr())
    assert L.b2c_conv3d_fwd(C.byref(desc), P(x), P(wp), P(y2), C.byref(fu2), P(ws), ws.numel(), gpu._st()) == 0
    torch.cuda.synchronize()
    assert torch.equal(st, st2)


def _nets(m):
    return nn.Sequential(m.ConvTranspose3d(32, 64, 4, 2, 1), m.BatchNorm3d(64), m.ReLU(True),
                         m.ConvTranspose3d(64, 160, 4, 2, 1), m.BatchNorm3d(160), m.ReLU(True),
                         m.Conv3d(160, 64, 4, 2, 1, bias=False), m.LayerNorm([64, 8, 8, 8]), m.LeakyReLU(0.2, inplace=True),
                         m.Conv3d(64, 32, 3, 1, 1), m.BatchNorm3d(32))


def check_module_level(dev, autocast):
    """fuse() wires conv -> BatchNorm3d / LayerNorm pairs to the epilogue statistics; results equal the unfused run."""
    import b200conv3d as B
    import b200conv3d.functional as BFn
    torch.manual_seed(0)
    net = _nets(B.nn).to(dev)
    B.fuse(net)
    assert [getattr(m, "fused_stats", None) for m in net if isinstance(m, (nn.Conv3d, nn.ConvTranspose3d))] == \
        ["channel", "channel", "sample", "channel"]
    x = torch.randn(4, 32, 4, 4, 4, device=dev)
    outs = []
    for on in (True, False):
        BFn.STATS[0] = on
        for p_ in net.parameters():
            p_.grad = None
        for m in net:
            if isinstance(m, nn.BatchNorm3d):
                m.reset_running_stats()
        used = []
        orig = BFn.take_stats
        BFn.take_stats = lambda *a: used.append(orig(*a)) or used[-1]
        try:
            ctx = torch.autocast("cuda", dtype=BF16) if autocast else torch.autocast("cpu", enabled=False)
            with ctx:
                y = net(x)
            y.float().square().mean().backward()
        finally:
            BFn.take_stats = orig
            BFn.STATS[0] = True
        assert [u is not None for u in used] == [on] * 4
        outs.append((y.detach().float(), [p_.grad.clone() for p_ in net.parameters()],
                     [m.running_var.clone() for m in net if isinstance(m, nn.BatchNorm3d)]))
    (ya, ga, ra), (yb, gb, rb) = outs
    tol = 2e-5
    assert ((ya - yb).abs().max() / yb.abs().max()).item() < (1e-2 if autocast else tol)
    for a, b in zip(ra, rb):
        assert ((a - b).abs().max() / b.abs().max()).item() < 1e-5
    gmax = max(g.abs().max().item() for g in gb)
    for a, b in zip(ga, gb):
        assert ((a - b).abs().max().item() / max(b.abs().max().item(), 1e-2 * gmax)) < (5e-2 if autocast else 1e-3)


@pytest.mark.gpu
@pytest.mark.parametrize("autocast", [False, True])
def test_gpu_module_level_fused_equals_unfused(autocast):
    check_module_level("cuda", autocast)


def test_host_module_level_fused_equals_unfused(oracle_path):
    code = '''
import sys
sys.path.insert(0, %r); sys.path.insert(0, %r)
from b200conv3d import _lib
_lib.use_library(%r, host=True)
import test_fused_stats as T
T.check_module_level("cpu", False)
print("HOST-OK")
''' % (os.path.join(ROOT, "3d-gan_b200"), os.path.join(ROOT, "tests"), oracle_path)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    assert "HOST-OK" in out.stdout, out.stdout[-2000:] + out.stderr[-4000:]
