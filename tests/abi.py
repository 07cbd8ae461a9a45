"""Test helper: drives ONE shared library that implements include/b2c.h on
tensors of one device.  The same calls run against the CPU oracle (host
pointers) and the product library (CUDA pointers), so parity tests are literal
A/B comparisons through the C ABI."""
import ctypes as C

import torch

from b200conv3d import _lib
from b200conv3d._lib import ConvDesc, Fusion

DT = {torch.bfloat16: 0, torch.float32: 2}


def make_desc(transposed, xshape, yshape, k, stride, pad, dil, dtype, algo=0):
    """xshape / yshape are logical (N, C, D, H, W)."""
    d = ConvDesc()
    d.x.n, d.x.c, d.x.d, d.x.h, d.x.w = xshape
    d.y.n, d.y.c, d.y.d, d.y.h, d.y.w = yshape
    for i in range(3):
        d.k[i], d.stride[i], d.pad[i], d.dil[i] = k[i], stride[i], pad[i], dil[i]
    d.transposed, d.io_dtype, d.algo = int(transposed), DT[dtype], algo
    return d


class Abi:
    def __init__(self, path, device):
        self.lib = _lib._bind(path)
        self.device = torch.device(device)
        self.cuda = self.device.type == "cuda"

    # -- plumbing
    def _p(self, t):
        return None if t is None else C.c_void_p(t.data_ptr())

    def _st(self):
        return C.c_void_p(torch.cuda.current_stream().cuda_stream) if self.cuda else None

    def _ck(self, rc, what):
        if rc != 0:
            raise RuntimeError("%s -> %d (%s)" % (what, rc, self.lib.b2c_strerror(rc).decode()))

    def _ws(self, n):
        return torch.empty(max(int(n), 16), dtype=torch.uint8, device=self.device)

    def to(self, t):
        return None if t is None else t.to(self.device).contiguous()

    # -- ops (all tensors NDHWC-dense, already on self.device)
    def pack(self, desc, pas, w):
        n = self.lib.b2c_filter_packed_bytes(C.byref(desc), pas)
        dt = torch.bfloat16 if desc.io_dtype == 0 else torch.float32
        out = torch.empty(n // (2 if desc.io_dtype == 0 else 4), dtype=dt, device=self.device)
        self._ck(self.lib.b2c_filter_pack(C.byref(desc), pas, self._p(w), self._p(out), self._st()), "pack")
        return out

    def conv_fwd(self, desc, x, w, bias=None, act=0, slope=0.0):
        wp = self.pack(desc, 0, w)
        dt = torch.bfloat16 if desc.io_dtype == 0 else torch.float32
        y = torch.empty((desc.y.n, desc.y.d, desc.y.h, desc.y.w, desc.y.c), dtype=dt, device=self.device)
        fu = Fusion(self._p(bias), act, slope)
        ws = self._ws(self.lib.b2c_conv3d_workspace(C.byref(desc), 0))
        self._ck(self.lib.b2c_conv3d_fwd(C.byref(desc), self._p(x), self._p(wp), self._p(y), C.byref(fu),
                                         self._p(ws), ws.numel(), self._st()), "conv_fwd")
        return y

    def conv_dgrad(self, desc, dy, w):
        wp = self.pack(desc, 1, w)
        dt = torch.bfloat16 if desc.io_dtype == 0 else torch.float32
        dx = torch.empty((desc.x.n, desc.x.d, desc.x.h, desc.x.w, desc.x.c), dtype=dt, device=self.device)
        ws = self._ws(self.lib.b2c_conv3d_workspace(C.byref(desc), 1))
        self._ck(self.lib.b2c_conv3d_dgrad(C.byref(desc), self._p(dy), self._p(wp), self._p(dx), self._p(ws),
                                           ws.numel(), self._st()), "conv_dgrad")
        return dx

    def conv_wgrad(self, desc, x, dy, wshape, want_bias=True):
        dw = torch.empty(wshape, dtype=torch.float32, device=self.device)
        db = torch.empty(desc.y.c, dtype=torch.float32, device=self.device) if want_bias else None
        ws = self._ws(self.lib.b2c_conv3d_workspace(C.byref(desc), 2))
        self._ck(self.lib.b2c_conv3d_wgrad(C.byref(desc), self._p(x), self._p(dy), self._p(dw), self._p(db),
                                           self._p(ws), ws.numel(), self._st()), "conv_wgrad")
        return dw, db

    def last_algo(self):
        return self.lib.b2c_last_algo()

    def sync(self):
        if self.cuda:
            torch.cuda.synchronize()


def rel_err(a, b):
    """max |a-b| / max |b| in fp64 on the CPU."""
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def kink_robust_err(a, b, tol, max_outlier_frac=0.01):
    """Parity criterion for bf16 gradients through piecewise-linear activations:
    an element whose pre-activation lies within bf16 rounding of the kink takes the
    other branch's slope (in stock torch autocast as well), so a small fraction of
    isolated elements may differ by O(slope gap * dy).  Returns the max error (relative to max|b|)
    over the non-outlier elements; raises if more than `max_outlier_frac` of the
    elements are outliers or the relative L2 error over ALL elements is >= 10*tol."""
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    scale = b.abs().max().clamp_min(1e-30)
    e = (a - b).abs() / scale
    out = e > tol
    frac = out.double().mean().item()
    assert frac <= max_outlier_frac, "outlier fraction %.4f" % frac
    l2 = ((a - b).norm() / b.norm().clamp_min(1e-30)).item()
    assert l2 < 10 * tol, "relative L2 error %.3e" % l2
    return e[~out].max().item() if (~out).any() else 0.0
