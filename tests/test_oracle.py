"""CPU: pins the plain-C oracle against torch's own CPU ops — the reference's
arithmetic lives in torch — through the same C ABI the product
library exports."""
import pytest
import torch
import torch.nn.functional as F

from abi import Abi, make_desc, rel_err

CONV_CASES = [
    # (transposed, cin, cout, k, stride, pad, dil, spatial)   reference site
    (False, 1, 8, 4, 2, 1, 1, (8, 8, 8)),      # dcgan.py:161 stem
    (False, 8, 16, 4, 2, 1, 1, (8, 8, 8)),     # dcgan.py:165
    (False, 16, 1, 4, 1, 0, 1, (4, 4, 4)),     # dcgan.py:181 head
    (False, 6, 6, 3, 1, 1, 1, (5, 6, 7)),      # utils.py:8-14 snconv3d 3^3
    (False, 6, 4, 1, 1, 0, 1, (4, 4, 4)),      # utils.py 1x1
    (False, 1, 4, 7, 2, 3, 1, (9, 9, 9)),      # FID_ResNet.py:119 stem
    (False, 4, 4, 3, 1, 2, 2, (7, 7, 7)),      # FID_ResNet.py layer3 dilation 2
    (False, 4, 4, 3, 1, 4, 4, (9, 9, 9)),      # FID_ResNet.py layer4 dilation 4
    (False, 4, 8, 1, 2, 0, 1, (6, 6, 6)),      # FID_ResNet.py:157 downsample
    (False, 4, 4, 3, 2, 0, 1, (9, 9, 9)),      # stylegan.py:257-260 down conv
    (True, 8, 16, 4, 1, 0, 1, (1, 1, 1)),      # dcgan.py:47
    (True, 8, 4, 4, 2, 1, 1, (4, 4, 4)),       # dcgan.py:51
    (True, 8, 1, 4, 2, 1, 1, (4, 4, 4)),       # dcgan.py:67
    (True, 4, 4, 3, 2, 0, 1, (5, 5, 5)),       # stylegan.py:278
]


@pytest.mark.parametrize("case", CONV_CASES)
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_conv_oracle_vs_torch(oracle_path, case, dtype):
    tr, cin, cout, k, s, p, d, sp = case
    torch.manual_seed(1)
    orc = Abi(oracle_path, "cpu")
    x = torch.randn(2, cin, *sp).to(dtype).float()
    w = (torch.randn(cin, cout, k, k, k) if tr else torch.randn(cout, cin, k, k, k)) * 0.2
    b = torch.randn(cout)
    wq = w.to(dtype).float()                       # what the packed filter holds
    xr = x.clone().requires_grad_(True)
    wr = wq.clone().requires_grad_(True)
    br = b.clone().requires_grad_(True)
    y = F.conv_transpose3d(xr, wr, br, s, p, 0, 1, d) if tr else F.conv3d(xr, wr, br, s, p, d)
    g = torch.randn_like(y).to(dtype).float()
    gx, gw, gb = torch.autograd.grad(y, (xr, wr, br), g)
    desc = make_desc(tr, tuple(x.shape), tuple(y.shape), (k,) * 3, (s,) * 3, (p,) * 3, (d,) * 3, dtype)
    nd = lambda t: t.permute(0, 2, 3, 4, 1).contiguous()
    tol = 1e-5 if dtype == torch.float32 else 8e-3
    yo = orc.conv_fwd(desc, nd(x).to(dtype), w, b)
    assert rel_err(yo, nd(y)) < tol
    dxo = orc.conv_dgrad(desc, nd(g).to(dtype), w)
    assert rel_err(dxo, nd(gx)) < tol
    dwo, dbo = orc.conv_wgrad(desc, nd(x).to(dtype), nd(g).to(dtype), w.shape)
    assert rel_err(dwo, gw) < 1e-5
    assert rel_err(dbo, gb) < 1e-5


def test_oracle_rejects_bad_desc(oracle_path):
    orc = Abi(oracle_path, "cpu")
    desc = make_desc(False, (1, 2, 4, 4, 4), (1, 2, 9, 9, 9), (3,) * 3, (1,) * 3, (1,) * 3, (1,) * 3, torch.float32)
    with pytest.raises(RuntimeError):
        orc.conv_fwd(desc, torch.zeros(1, 4, 4, 4, 2), torch.zeros(2, 2, 3, 3, 3))
