"""GPU parity: product library (CUDA pointers) vs the CPU oracle (host pointers)
through the same C ABI on the same seeded inputs.  Tolerances are north_star's:
fp32 max-rel <= 1e-4, bf16 <= 2e-2 (relative to max|ref| of the tensor)."""
import ctypes as C

import pytest
import torch

from abi import Abi, make_desc, rel_err

pytestmark = pytest.mark.gpu
TOL = {torch.float32: 1e-4, torch.bfloat16: 2e-2}
AUTO, DIRECT, TC = 0, 1, 2

# (transposed, cin, cout, k, stride, pad, dil, spatial, batch)    reference site
TC_CASES = [
    (False, 64, 128, 4, 2, 1, 1, (8, 8, 8), 2),     # dcgan.py:165-177 D body
    (False, 128, 64, 4, 2, 1, 1, (16, 16, 16), 3),  # several M tiles, 8 phase maps
    (False, 64, 64, 3, 1, 1, 1, (5, 6, 7), 2),      # utils.py snconv3d 3^3, ragged tiles
    (False, 64, 64, 3, 1, 2, 2, (8, 8, 8), 2),      # FID_ResNet.py layer3 (dilation 2)
    (False, 64, 64, 3, 1, 4, 4, (8, 8, 8), 1),      # FID_ResNet.py layer4 (dilation 4)
    (False, 64, 256, 1, 1, 0, 1, (8, 8, 8), 2),     # FID_ResNet.py:79 1x1 expand
    (False, 128, 64, 1, 2, 0, 1, (8, 8, 8), 2),     # FID_ResNet.py:157 1x1 stride 2
    (False, 64, 64, 3, 2, 0, 1, (9, 9, 9), 2),      # stylegan.py:257-260 down conv
    (False, 32, 32, 3, 1, 1, 1, (8, 8, 8), 2),      # biggan hidden 32 ch (K tail < 64)
    (False, 96, 24, 3, 1, 1, 1, (6, 6, 6), 2),      # K tail (96 = 64+32), N not /16
    (False, 64, 320, 1, 1, 0, 1, (4, 4, 4), 2),     # two N tiles (256 + 64)
    (False, 256, 16, 4, 1, 0, 1, (4, 4, 4), 4),     # dcgan.py:181-like head (valid conv)
    (True, 128, 64, 4, 2, 1, 1, (4, 4, 4), 2),      # dcgan.py:51-65 G body (sub-pixel)
    (True, 64, 64, 4, 2, 1, 1, (8, 8, 8), 3),
    (True, 64, 128, 4, 1, 0, 1, (1, 1, 1), 4),      # dcgan.py:47 z -> 4^3 (point GEMM)
    (True, 64, 32, 3, 2, 0, 1, (5, 5, 5), 2),       # stylegan.py:278 up conv (odd output)
    (True, 64, 64, 3, 1, 1, 1, (5, 5, 5), 2),
]
DIRECT_CASES = [
    (False, 1, 16, 4, 2, 1, 1, (16, 16, 16), 2),    # dcgan.py:161 stem (Cin = 1)
    (False, 64, 1, 4, 1, 0, 1, (4, 4, 4), 4),       # dcgan.py:181 head (Cout = 1)
    (False, 1, 8, 7, 2, 3, 1, (12, 12, 12), 2),     # FID_ResNet.py:119 stem
    (False, 6, 10, 3, 1, 1, 1, (5, 6, 7), 2),
    (False, 40, 40, 3, 1, 2, 2, (7, 7, 7), 2),
    (True, 16, 1, 4, 2, 1, 1, (8, 8, 8), 2),        # dcgan.py:67 G head (Cout = 1)
    (True, 8, 12, 4, 1, 0, 1, (1, 1, 1), 3),
    (True, 40, 8, 3, 2, 0, 1, (5, 5, 5), 2),
]
# big-side tensor has ONE channel: im2col / col2im + tcgen05 GEMM (csrc/thin.cu)
THIN_CASES = [
    (False, 1, 16, 4, 2, 1, 1, (16, 16, 16), 2),    # dcgan.py:161 stem
    (False, 1, 64, 4, 2, 1, 1, (8, 10, 12), 3),     # ragged spatial extents
    (False, 1, 32, 7, 2, 3, 1, (12, 12, 12), 2),    # FID_ResNet.py:119 stem (Kp = 344)
    (False, 1, 64, 3, 1, 1, 1, (6, 7, 8), 2),       # biggan.py:78 input conv (Kp = 32)
    (False, 1, 64, 3, 1, 1, 1, (4, 5, 32), 2),      # same, W % 32 == 0: the k3 / s1 line loader (shuffled neighbours)
    (False, 1, 32, 3, 1, 1, 1, (3, 3, 64), 3),      # two warps per W line
    (False, 64, 1, 3, 1, 1, 1, (4, 5, 32), 2),      # biggan.py:42 output conv: its data gradient takes the line loader
    (True, 16, 1, 4, 2, 1, 1, (8, 8, 8), 2),        # dcgan.py:67 G head
    (True, 128, 1, 4, 2, 1, 1, (6, 6, 6), 2),       # dcgan.py:67 at full width
    (True, 16, 1, 4, 2, 1, 1, (20, 9, 70), 2),      # TMA-staged col2im: three d chunks, ragged h tiles, two w tiles
    (False, 1, 16, 4, 2, 1, 1, (18, 10, 132), 1),   # its data-gradient twin (dcgan.py:161), small side 9 x 5 x 66
    # SMALL side has one channel, stride 1: mirrored problem (thin_mode == 2)
    (False, 64, 1, 3, 1, 1, 1, (6, 7, 8), 2),       # biggan.py:42 output conv f -> 1
    (False, 16, 1, 3, 1, 1, 1, (9, 8, 10), 3),      # stylegan.py:1038 Conv3d(16 -> 1)
    (False, 64, 1, 4, 1, 0, 1, (4, 4, 4), 4),       # dcgan.py:181 head (valid conv, mirrored pad 3)
    (False, 32, 1, 3, 1, 2, 2, (8, 8, 8), 2),       # dilated
    (False, 24, 1, 1, 1, 0, 1, (5, 6, 7), 2),       # 1x1x1 to one channel (Kp = 16)
]


@pytest.fixture(scope="module")
def libs(product_path, oracle_path):
    gpu, cpu = Abi(product_path, "cuda:0"), Abi(oracle_path, "cpu")
    dbg = C.CDLL(product_path)
    dbg.b2c_debug_tc_error.restype = C.c_uint
    dbg.b2c_debug_tc_error()          # clear
    return gpu, cpu, dbg


def _run(libs, case, dtype, algo, act=0, bias=True):
    gpu, cpu, dbg = libs
    tr, cin, cout, k, s, p, d, sp, n = case
    torch.manual_seed(hash(case) % 1000)
    osp = [((v - 1) * s - 2 * p + d * (k - 1) + 1) if tr else ((v + 2 * p - d * (k - 1) - 1) // s + 1) for v in sp]
    x = torch.randn(n, *sp, cin).to(dtype)
    dy = torch.randn(n, *osp, cout).to(dtype)
    w = (torch.randn(cin, cout, k, k, k) if tr else torch.randn(cout, cin, k, k, k)) / (cin * k ** 3) ** 0.5
    b = torch.randn(cout) if bias else None
    desc = make_desc(tr, (n, cin, *sp), (n, cout, *osp), (k,) * 3, (s,) * 3, (p,) * 3, (d,) * 3, dtype, algo)
    dref = make_desc(tr, (n, cin, *sp), (n, cout, *osp), (k,) * 3, (s,) * 3, (p,) * 3, (d,) * 3, dtype, 0)
    tol = TOL[dtype]
    errs = {}
    y = gpu.conv_fwd(desc, gpu.to(x), gpu.to(w), gpu.to(b), act, 0.2)
    used = gpu.last_algo()
    errs["fwd"] = rel_err(y, cpu.conv_fwd(dref, x, w, b, act, 0.2))
    dx = gpu.conv_dgrad(desc, gpu.to(dy), gpu.to(w))
    errs["dgrad"] = rel_err(dx, cpu.conv_dgrad(dref, dy, w))
    dw, db = gpu.conv_wgrad(desc, gpu.to(x), gpu.to(dy), w.shape)
    dwr, dbr = cpu.conv_wgrad(dref, x, dy, w.shape)
    errs["wgrad"] = rel_err(dw, dwr)
    errs["dbias"] = rel_err(db, dbr)
    flag = dbg.b2c_debug_tc_error()
    assert flag == 0, "tensor-core kernel reported error word 0x%x (%s)" % (flag, errs)
    if algo != AUTO:
        assert used == algo
    bad = {k_: v for k_, v in errs.items() if not v < (1e-4 if k_ in ("wgrad", "dbias") and dtype == torch.float32 else tol)}
    assert not bad, "case %s dtype %s algo %d: %s" % (case, dtype, algo, errs)
    return errs


@pytest.mark.parametrize("case", TC_CASES)
def test_tensor_core_path_bf16(libs, case):
    _run(libs, case, torch.bfloat16, TC)


@pytest.mark.parametrize("case", TC_CASES[:4] + DIRECT_CASES)
def test_direct_path_bf16(libs, case):
    _run(libs, case, torch.bfloat16, DIRECT)


@pytest.mark.parametrize("case", TC_CASES[:3] + TC_CASES[12:14] + DIRECT_CASES)
def test_direct_path_fp32(libs, case):
    _run(libs, case, torch.float32, AUTO)


@pytest.mark.parametrize("case", THIN_CASES)
def test_thin_path_bf16(libs, case):
    _run(libs, case, torch.bfloat16, AUTO)
    assert libs[0].last_algo() == TC
    _run(libs, case, torch.bfloat16, AUTO, act=3)      # tanh epilogue (dcgan.py:68)


@pytest.mark.parametrize("act", [1, 2, 3])
def test_fused_activation_epilogue(libs, act):
    _run(libs, TC_CASES[0], torch.bfloat16, TC, act=act)
    _run(libs, DIRECT_CASES[0], torch.bfloat16, AUTO, act=act)


def test_auto_prefers_tensor_cores_for_bf16(libs):
    gpu = libs[0]
    _run(libs, TC_CASES[12], torch.bfloat16, AUTO)
    assert gpu.last_algo() == TC


def test_linearity_full_width_layer(libs):
    """Size-independent property at a BASELINE-sized layer (dcgan.py:173,
    512->1024 ch, 16^3->8^3, batch 4): conv(a*x1 + x2) == a*conv(x1) + conv(x2)."""
    gpu = libs[0]
    torch.manual_seed(5)
    n, cin, cout, sp = 4, 512, 1024, 16
    desc = make_desc(False, (n, cin, sp, sp, sp), (n, cout, 8, 8, 8), (4,) * 3, (2,) * 3, (1,) * 3, (1,) * 3,
                     torch.bfloat16, TC)
    w = torch.randn(cout, cin, 4, 4, 4, device="cuda") * 0.01
    x1 = torch.randn(n, sp, sp, sp, cin, device="cuda").to(torch.bfloat16)
    x2 = torch.randn(n, sp, sp, sp, cin, device="cuda").to(torch.bfloat16)
    y1, y2 = gpu.conv_fwd(desc, x1, w).float(), gpu.conv_fwd(desc, x2, w).float()
    y12 = gpu.conv_fwd(desc, (2 * x1.float() + x2.float()).to(torch.bfloat16), w).float()
    assert libs[2].b2c_debug_tc_error() == 0
    assert rel_err(y12, 2 * y1 + y2) < 2e-2
