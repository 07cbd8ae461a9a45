"""W-window lowering of the narrow 3-tap convolutions (include/b2c.h b2c_window_w; functional._conv3d_windowed):
utils.py:57-60,105-109 hidden 32- / 64-channel 3^3 convs, stylegan.py:285-287.

CPU: the oracle's window gather against torch's own unfold of the W axis (forward + adjoint), and the windowed
convolution through the oracle library against F.conv3d (forward, dx, dw, dbias).  GPU: the CUDA gather against the
oracle bit for bit, the windowed convolution on the tcgen05 path against F.conv3d, and equality with the folded path."""
import os
import subprocess
import sys

import pytest
import torch
import torch.nn.functional as F

from abi import Abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DT = {torch.bfloat16: 0, torch.float32: 2}
CASES = [(6, 8, 16, 4), (5, 12, 8, 2), (3, 4, 24, 4), (2, 6, 5, 1), (4, 16, 3, 8)]      # rows, W, C, F


def _window_ref(x, f):
    """torch: pad W by one voxel each side, unfold windows of F + 2 voxels at stride F."""
    rows, w, c = x.shape
    xp = F.pad(x, (0, 0, 1, 1))
    return xp.unfold(1, f + 2, f).permute(0, 1, 3, 2).reshape(rows, w // f, (f + 2) * c)


def _run(abi, x, f, dtype, inverse_of=None):
    rows, w, c = x.shape
    src = abi.to((inverse_of if inverse_of is not None else x).to(dtype))
    shape = (rows, w, c) if inverse_of is not None else (rows, w // f, (f + 2) * c)
    dst = torch.empty(shape, dtype=dtype, device=abi.device)
    abi._ck(abi.lib.b2c_window_w(abi._p(src), abi._p(dst), rows, w, c, f, 1, DT[dtype], 1 if inverse_of is not None else 0,
                                 abi._st()), "window_w")
    return dst.float().cpu()


@pytest.mark.parametrize("case", CASES)
def test_oracle_window_matches_torch_unfold(oracle_path, case):
    rows, w, c, f = case
    cpu = Abi(oracle_path, "cpu")
    g = torch.Generator().manual_seed(sum(case))
    x = torch.randn(rows, w, c, generator=g)
    y = _run(cpu, x, f, torch.float32)
    xt = x.clone().requires_grad_(True)
    yt = _window_ref(xt, f)
    torch.testing.assert_close(y, yt.detach(), rtol=0, atol=0)
    gy = torch.randn(yt.shape, generator=g)
    yt.backward(gy)
    dx = _run(cpu, x, f, torch.float32, inverse_of=gy)
    torch.testing.assert_close(dx, xt.grad, rtol=1e-6, atol=1e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("case", CASES)
def test_gpu_window_matches_oracle(product_path, oracle_path, case, dtype):
    rows, w, c, f = case
    gpu, cpu = Abi(product_path, "cuda:0"), Abi(oracle_path, "cpu")
    g = torch.Generator().manual_seed(sum(case))
    x = torch.randn(rows, w, c, generator=g).to(dtype).float()
    torch.testing.assert_close(_run(gpu, x, f, dtype), _run(cpu, x, f, dtype), rtol=0, atol=0)
    gy = torch.randn(rows, w // f, (f + 2) * c, generator=g).to(dtype).float()
    tol = 1e-6 if dtype == torch.float32 else 1e-2      # up to F+2 terms per voxel, summed in another order
    torch.testing.assert_close(_run(gpu, x, f, dtype, inverse_of=gy), _run(cpu, x, f, dtype, inverse_of=gy), rtol=tol, atol=tol)


CONV_CASES = [(32, 32, (3, 4, 8)), (64, 64, (2, 3, 6)), (16, 32, (2, 2, 12)), (48, 64, (3, 2, 4))]     # cin, cout, (d, h, w)


def check_windowed_conv(dev):
    """conv3d(k3, s1, p1) with 32 / 64 output channels takes the window lowering and equals F.conv3d."""
    import b200conv3d.functional as BF

    def rel(a, b):
        a, b = a.double().cpu(), b.double().cpu()
        return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()

    torch.manual_seed(0)
    for cin, cout, sp in CONV_CASES:
        x = torch.randn(2, cin, *sp).to(torch.bfloat16).to(dev).requires_grad_(True)
        w = (torch.randn(cout, cin, 3, 3, 3) * 0.1).to(dev).requires_grad_(True)
        b = torch.randn(cout).to(dev).requires_grad_(True)
        assert BF._window_factor(x, w, BF.Geom(False, (3, 3, 3), 1, 1, 1)) == 128 // cout
        y = BF.conv3d(x, w, b, 1, 1, act="lrelu", slope=0.2)
        assert tuple(y.shape) == (2, cout, *sp)
        xr = x.detach().float().cpu().requires_grad_(True)
        wr = w.detach().cpu().to(torch.bfloat16).float().requires_grad_(True)
        br = b.detach().cpu().clone().requires_grad_(True)
        yr = F.leaky_relu(F.conv3d(xr, wr, br, 1, 1), 0.2)
        gy = torch.randn_like(yr).to(torch.bfloat16)
        y.backward(gy.to(dev))
        yr.backward(gy.float())
        e = dict(fwd=rel(y.float(), yr), dx=rel(x.grad.float(), xr.grad), dw=rel(w.grad, wr.grad), db=rel(b.grad, br.grad))
        assert all(v < 2e-2 for v in e.values()), (cin, cout, sp, e)
        # the older three-tap folding gives the same result (different summation order only)
        BF.WINDOW[0] = False
        try:
            y2 = BF.conv3d(x.detach(), w.detach(), b.detach(), 1, 1, act="lrelu", slope=0.2)
        finally:
            BF.WINDOW[0] = True
        assert rel(y2.float(), y.detach().float()) < 1e-2


def test_host_windowed_conv_equals_torch(oracle_path):
    code = '''
import sys
sys.path.insert(0, %r); sys.path.insert(0, %r)
from b200conv3d import _lib
_lib.use_library(%r, host=True)
import test_window_conv as T
T.check_windowed_conv("cpu")
print("HOST-OK")
''' % (os.path.join(ROOT, "3d-gan_b200"), os.path.join(ROOT, "tests"), oracle_path)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    assert "HOST-OK" in out.stdout, out.stdout[-2000:] + out.stderr[-4000:]


@pytest.mark.gpu
def test_gpu_windowed_conv_equals_torch():
    check_windowed_conv("cuda")
    from b200conv3d import _lib
    assert _lib.lib().b2c_last_algo() == 2          # the re-shaped problem ran on the tcgen05 path
