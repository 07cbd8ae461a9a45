"""x2 resampling ABI (include/b2c.h b2c_upsample2x_* / b2c_avgpool2_*).

CPU part: the C oracle against the reference's own calls - `x[:, :C]` + F.interpolate(
scale_factor=2) (utils.py:76-80) and nn.AvgPool3d(2) (utils.py:113-116) - forward and autograd
backward, in torch on the CPU.  GPU part: the CUDA kernels against the oracle through the C ABI
(bit-exact in fp32: pure copies / 8-term sums; bf16 within one rounding), and the autograd
Functions against torch."""
import ctypes as C

import pytest
import torch
import torch.nn.functional as F

from abi import Abi

DT = {torch.bfloat16: 0, torch.float32: 2}
CASES = [  # n, d, h, w, cx, c
    (2, 3, 4, 5, 16, 16), (2, 4, 4, 4, 32, 8), (1, 2, 3, 2, 6, 5), (3, 1, 2, 8, 24, 24),
]


def _run_abi(abi, dtype, case):
    n, d, h, w, cx, c = case
    g = torch.Generator().manual_seed(sum(case))
    x = torch.randn(n, d, h, w, cx, generator=g).to(dtype)
    dy = torch.randn(n, 2 * d, 2 * h, 2 * w, c, generator=g).to(dtype)
    xd, dyd = abi.to(x), abi.to(dy)
    y = torch.empty(n, 2 * d, 2 * h, 2 * w, c, dtype=dtype, device=abi.device)
    dx = torch.empty(n, d, h, w, cx, dtype=dtype, device=abi.device)
    st, P = abi._st(), abi._p
    abi._ck(abi.lib.b2c_upsample2x_fwd(P(xd), P(y), n, d, h, w, cx, 0, c, DT[dtype], st), "up fwd")
    abi._ck(abi.lib.b2c_upsample2x_bwd(P(dyd), P(dx), n, d, h, w, cx, 0, c, DT[dtype], st), "up bwd")
    # avg pool on the up-sampled-size tensor dy: [n, 2d, 2h, 2w, c] -> [n, d, h, w, c]
    py = torch.empty(n, d, h, w, c, dtype=dtype, device=abi.device)
    pdx = torch.empty_like(dyd)
    pg = abi.to(torch.randn(n, d, h, w, c, generator=g).to(dtype))
    abi._ck(abi.lib.b2c_avgpool2_fwd(P(dyd), P(py), n, 2 * d, 2 * h, 2 * w, c, DT[dtype], st), "pool fwd")
    abi._ck(abi.lib.b2c_avgpool2_bwd(P(pg), P(pdx), n, 2 * d, 2 * h, 2 * w, c, DT[dtype], st), "pool bwd")
    return x, dy, pg.cpu(), [t.float().cpu() for t in (y, dx, py, pdx)]


@pytest.mark.parametrize("case", CASES)
def test_oracle_resample_matches_torch(oracle_path, case):
    cpu = Abi(oracle_path, "cpu")
    n, d, h, w, cx, c = case
    x, dy, pg, (y, dx, py, pdx) = _run_abi(cpu, torch.float32, case)
    xt = x.permute(0, 4, 1, 2, 3).clone().requires_grad_(True)
    yt = F.interpolate(xt[:, :c], scale_factor=2)
    yt.backward(dy.permute(0, 4, 1, 2, 3))
    torch.testing.assert_close(y, yt.detach().permute(0, 2, 3, 4, 1), rtol=0, atol=0)
    torch.testing.assert_close(dx, xt.grad.permute(0, 2, 3, 4, 1), rtol=1e-6, atol=1e-6)
    pt = dy.permute(0, 4, 1, 2, 3).clone().requires_grad_(True)
    pyt = torch.nn.AvgPool3d(2)(pt)
    pyt.backward(pg.permute(0, 4, 1, 2, 3))
    torch.testing.assert_close(py, pyt.detach().permute(0, 2, 3, 4, 1), rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(pdx, pt.grad.permute(0, 2, 3, 4, 1), rtol=1e-6, atol=1e-7)


def test_oracle_avgpool_floor_on_odd_extents(oracle_path):
    cpu = Abi(oracle_path, "cpu")
    x = torch.randn(2, 5, 4, 7, 3)
    y = torch.empty(2, 2, 2, 3, 3)
    cpu._ck(cpu.lib.b2c_avgpool2_fwd(cpu._p(x), cpu._p(y), 2, 5, 4, 7, 3, 2, None), "pool")
    ref = torch.nn.AvgPool3d(2)(x.permute(0, 4, 1, 2, 3)).permute(0, 2, 3, 4, 1)
    torch.testing.assert_close(y, ref, rtol=1e-6, atol=1e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("case", CASES)
def test_gpu_resample_matches_oracle(product_path, oracle_path, case, dtype):
    gpu, cpu = Abi(product_path, "cuda:0"), Abi(oracle_path, "cpu")
    _, _, _, got = _run_abi(gpu, dtype, case)
    _, _, _, want = _run_abi(cpu, dtype, case)
    tol = 1e-6 if dtype == torch.float32 else 1e-2     # fp32: sums of 8 in fp32 vs the oracle's double
    for a, b, name in zip(got, want, ("up_fwd", "up_bwd", "pool_fwd", "pool_bwd")):
        torch.testing.assert_close(a, b, rtol=tol, atol=tol * 4, msg=lambda m: name + ": " + m)


@pytest.mark.gpu
def test_gpu_resample_autograd_vs_torch():
    import b200conv3d as B
    torch.manual_seed(3)
    x = torch.randn(2, 16, 3, 4, 5)
    xr = x.clone().requires_grad_(True)
    xo = x.cuda().requires_grad_(True)
    yr = torch.nn.AvgPool3d(2)(F.interpolate(xr[:, :8], scale_factor=2) * 1.5)
    yo = B.nn.AvgPool3d(2)(B.functional.upsample_nearest2x(xo, 8) * 1.5)
    g = torch.randn_like(yr)
    yr.backward(g), yo.backward(g.cuda())
    torch.testing.assert_close(yo.cpu(), yr, rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(xo.grad.cpu(), xr.grad, rtol=1e-6, atol=1e-6)
    with pytest.raises(ValueError):
        B.nn.AvgPool3d(3)(xo)


def _maxpool_cases():
    return [((2, 8, 9, 10, 11), 3, 2, 1), ((1, 16, 8, 8, 8), 3, 2, 1), ((2, 3, 7, 6, 5), 2, 2, 0), ((1, 8, 6, 6, 6), 3, 1, 1)]


def check_maxpool(dev, dtype):
    """b2c_maxpool3d fwd / bwd against F.max_pool3d (FID_ResNet.py:129 uses k3 s2 p1)."""
    import torch
    import torch.nn.functional as F
    import b200conv3d.functional as BF
    for shape, k, s, p in _maxpool_cases():
        g = torch.Generator().manual_seed(sum(shape))
        x = torch.randn(*shape, generator=g).to(dtype).float()
        x[0, 0, :2, :2, :2] = 1.5                      # ties: the gradient must go to the first maximum only
        xr = x.clone().requires_grad_(True)
        yr = F.max_pool3d(xr, k, s, p)
        gy = torch.randn(*yr.shape, generator=g).to(dtype).float()
        yr.backward(gy)
        xg = x.clone().to(dev).to(dtype).requires_grad_(True)
        y = BF.max_pool3d(xg, k, s, p)
        y.backward(gy.to(dev).to(dtype))
        assert y.shape == yr.shape
        assert torch.equal(y.float().cpu(), yr.detach()), (shape, k, s, p)
        tol = 1e-6 if dtype == torch.float32 else 1e-2
        assert (xg.grad.float().cpu() - xr.grad).abs().max().item() <= tol * xr.grad.abs().max().item(), (shape, k, s, p)


def test_maxpool_oracle_library_vs_torch(oracle_path):
    import os
    import subprocess
    import sys
    ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = r'''
import sys, torch
sys.path.insert(0, %r); sys.path.insert(0, %r)
from b200conv3d import _lib
_lib.use_library(%r, host=True)
import test_resample as T
T.check_maxpool("cpu", torch.float32)
print("HOST-OK")
''' % (os.path.join(ROOT, "3d-gan_b200"), os.path.join(ROOT, "tests"), oracle_path)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    assert "HOST-OK" in out.stdout, out.stdout[-2000:] + out.stderr[-4000:]


@pytest.mark.gpu
def test_gpu_maxpool_vs_torch():
    import torch
    check_maxpool("cuda", torch.float32)
    check_maxpool("cuda", torch.bfloat16)


def check_d2s(dev, dtype):
    """b2c_depth_to_space2 (both directions, cropped and even grids) against the view / permute formulation."""
    import torch
    import b200conv3d.functional as BF
    for n, c, j, crop in ((2, 8, (3, 4, 5), (1, 1, 1)), (1, 16, (2, 2, 3), (0, 0, 0)), (2, 3, (2, 3, 2), (1, 0, 1)),
                          (1, 32, (4, 1, 2), (0, 1, 0))):
        g = torch.Generator().manual_seed(n + c + sum(j))
        out = tuple(2 * a - b for a, b in zip(j, crop))
        z = torch.randn(n, 8 * c, *j, generator=g).to(dtype).float()
        zr = z.clone().requires_grad_(True)
        yr = zr.reshape(n, 2, 2, 2, c, *j).permute(0, 4, 5, 1, 6, 2, 7, 3).reshape(n, c, 2 * j[0], 2 * j[1], 2 * j[2])
        yr = yr[:, :, :out[0], :out[1], :out[2]]
        gy = torch.randn(*yr.shape, generator=g).to(dtype).float()
        yr.backward(gy)
        zg = z.clone().to(dev).to(dtype).requires_grad_(True)
        y = BF._D2SFn.apply(zg, out)
        y.backward(gy.to(dev).to(dtype))
        assert y.shape == yr.shape
        assert torch.equal(y.float().cpu(), yr.detach()), (n, c, j, crop)
        assert torch.equal(zg.grad.float().cpu(), zr.grad), (n, c, j, crop)


def test_d2s_oracle_library_vs_torch(oracle_path):
    import os
    import subprocess
    import sys
    ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = r'''
import sys, torch
sys.path.insert(0, %r); sys.path.insert(0, %r)
from b200conv3d import _lib
_lib.use_library(%r, host=True)
import test_resample as T
T.check_d2s("cpu", torch.float32)
print("HOST-OK")
''' % (os.path.join(ROOT, "3d-gan_b200"), os.path.join(ROOT, "tests"), oracle_path)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    assert "HOST-OK" in out.stdout, out.stdout[-2000:] + out.stderr[-4000:]


@pytest.mark.gpu
def test_gpu_d2s_vs_torch():
    import torch
    check_d2s("cuda", torch.float32)
    check_d2s("cuda", torch.bfloat16)
