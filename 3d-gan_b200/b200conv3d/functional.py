"""autograd.Functions over the C ABI (include/b2c.h).

Mirrors the torch functional API the reference calls (SURVEY.md §8(b) Face 1):
``F.conv3d`` / ``F.conv_transpose3d`` (stylegan.py:168-170), and the fused
BatchNorm3d / LayerNorm / activation epilogues of dcgan.py:45-70,158-182,
utils.py:47-132, FID_ResNet.py:65-102.

Tensors cross this layer as logical NCDHW of any stride; internally everything
is dense NDHWC (torch ``channels_last_3d``), bf16 under autocast / bf16 inputs
and fp32 otherwise.  Every FLOP runs in libb2c.so; there is no torch/cuDNN
compute fallback.
"""
import ctypes as C

import torch
from torch.autograd import Function

from . import _lib
from ._lib import (ACT_IDS, ACT_NONE, ALGO_AUTO, BF16, F32, PASS_DGRAD,
                   PASS_FWD, PASS_WGRAD, STAT_CHANNEL, STAT_SAMPLE, ConvDesc, Fusion, check)

_force_algo = [ALGO_AUTO]     # tests / profiling: force DIRECT or TC
PROFILE = None                # bench.py: list collecting (pass, desc tuple, start event, end event)


class _timed:
    """Bracket one C-ABI conv launch with CUDA events on its stream when bench.py has
    switched per-launch profiling on (PROFILE is a list)."""

    def __init__(self, pas, desc, fold=1, orig=None):
        self.on = PROFILE is not None
        if self.on:       # W-folded launches are reported with the ORIGINAL channel counts (useful FLOPs)
            xc, yc = orig if orig is not None else (desc.x.c // fold, desc.y.c // fold)
            self.key = (pas, bool(desc.transposed), desc.x.n, xc, desc.x.d, yc, desc.y.d,
                        desc.k[0], desc.stride[0], desc.pad[0], desc.dil[0])

    def __enter__(self):
        if self.on:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e1 = torch.cuda.Event(enable_timing=True)
            self.e0.record()

    def __exit__(self, *a):
        if self.on:
            self.e1.record()
            PROFILE.append((self.key, self.e0, self.e1, _lib.lib().b2c_last_algo()))



def set_algo(algo):
    _force_algo[0] = int(algo)


# ----------------------------------------------------------------- helpers
def _dt(dtype):
    if dtype == torch.bfloat16:
        return BF16
    if dtype == torch.float32:
        return F32
    raise TypeError("b200conv3d supports bf16 and fp32 compute, got %s" % dtype)


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream(t):
    if _lib.is_host():
        return None
    return C.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


def _check_device(t):
    host = _lib.is_host()
    if host and t.is_cuda:
        raise RuntimeError("b200conv3d is bound to the host oracle; got a CUDA tensor")
    if not host and not t.is_cuda:
        raise RuntimeError(
            "b200conv3d ops run only on CUDA (sm_100a) tensors; got a %s tensor — "
            "there is no CPU fallback" % t.device)


class _dev:
    """Make the tensor's device current while launching (DataParallel threads)."""

    def __init__(self, t):
        self.g = None if _lib.is_host() else torch.cuda.device(t.device)

    def __enter__(self):
        if self.g is not None:
            self.g.__enter__()

    def __exit__(self, *a):
        if self.g is not None:
            self.g.__exit__(*a)


def compute_dtype(x):
    """bf16 under any autocast (the reference asks for fp16 autocast,
    trainer.py:213; the B200 path computes in bf16 — BASELINE config 2) or for
    bf16 inputs; fp32 otherwise."""
    kind = "cpu" if _lib.is_host() else "cuda"
    if torch.is_autocast_enabled(kind):
        return torch.bfloat16
    if x.dtype in (torch.bfloat16, torch.float16):
        return torch.bfloat16
    return torch.float32


def _ws(nbytes, like):
    return torch.empty(max(int(nbytes), 16), dtype=torch.uint8, device=like.device)


def cast(x, dtype):
    """Dense elementwise dtype conversion through b2c_cast."""
    if x.dtype == dtype:
        return x
    if x.dtype not in (torch.bfloat16, torch.float32):
        x = x.float()                      # fp16/fp64 plumbing only
    assert x.is_contiguous()
    out = torch.empty_like(x, dtype=dtype)
    with _dev(x):
        check(_lib.lib().b2c_cast(_ptr(x), _ptr(out), x.numel(), _dt(x.dtype), _dt(dtype),
                                  _stream(x)), "cast")
    return out


def to_ndhwc(x, dtype):
    """logical [N,C,D,H,W] (any stride/dtype) -> dense [N,D,H,W,C] of `dtype`."""
    _check_device(x)
    if x.dim() != 5:
        raise ValueError("expected a 5-D NCDHW tensor, got shape %s" % (tuple(x.shape),))
    p = x.permute(0, 2, 3, 4, 1)
    if p.is_contiguous():
        return p if x.dtype == dtype else cast(p, dtype)
    if x.dtype not in (torch.bfloat16, torch.float32):
        x = x.float()
        p = x.permute(0, 2, 3, 4, 1)
    if p.stride(4) == 1 and x.shape[1] > 1 and not x.is_contiguous():
        # a channel slice / crop of NDHWC storage (`x[:, :Cout]`, utils.py:74): one strided copy of whole channel
        # runs instead of the NCDHW round trip
        pc = p.contiguous()
        return pc if pc.dtype == dtype else cast(pc, dtype)
    xs = x if x.is_contiguous() else x.contiguous()
    n, c, d, h, w = xs.shape
    out = torch.empty((n, d, h, w, c), dtype=dtype, device=x.device)
    with _dev(x):
        check(_lib.lib().b2c_transpose(_ptr(xs), _ptr(out), n, c, d * h * w, _dt(xs.dtype),
                                       _dt(dtype), _stream(x)), "transpose")
    return out


def from_ndhwc(buf):
    """dense [N,D,H,W,C] -> logical NCDHW view (channels_last_3d strides)."""
    return buf.permute(0, 4, 1, 2, 3)


def alloc_out(n, d, h, w, c, dtype, device):
    """Fresh logical-NCDHW tensor whose storage is dense NDHWC.  Not a view, so the
    reference's in-place activations (nn.ReLU(True), dcgan.py:49) may modify the
    output of a custom Function.  Returns (logical tensor, NDHWC view of it)."""
    t = torch.empty_strided((n, c, d, h, w), (d * h * w * c, 1, h * w * c, w * c, c),
                            dtype=dtype, device=device)
    return t, t.permute(0, 2, 3, 4, 1)


def _triple(v):
    if isinstance(v, (list, tuple)):
        if len(v) == 1:
            return (int(v[0]),) * 3
        if len(v) != 3:
            raise ValueError("expected an int or 3 ints, got %r" % (v,))
        return tuple(int(i) for i in v)
    return (int(v),) * 3


class Geom:
    """Static geometry of one conv / conv-transpose call."""
    __slots__ = ("transposed", "k", "stride", "pad", "dil", "out_pad", "fold", "orig")

    def __init__(self, transposed, k, stride, pad, dil, out_pad=(0, 0, 0)):
        self.fold = 1                 # W-folding factor of the launch (output channels = fold x the layer's)
        self.orig = None              # (Cin, Cout) of the layer when the launch runs a re-shaped problem (profiling)
        self.transposed, self.k = bool(transposed), tuple(k)
        self.stride, self.pad, self.dil = _triple(stride), _triple(pad), _triple(dil)
        self.out_pad = _triple(out_pad)

    def out_spatial(self, sp):
        o = []
        for i in range(3):
            if not self.transposed:
                v = (sp[i] + 2 * self.pad[i] - self.dil[i] * (self.k[i] - 1) - 1) // self.stride[i] + 1
            else:
                v = ((sp[i] - 1) * self.stride[i] - 2 * self.pad[i]
                     + self.dil[i] * (self.k[i] - 1) + self.out_pad[i] + 1)
            if v <= 0:
                raise RuntimeError("conv3d: computed output size %d is too small" % v)
            o.append(v)
        return o


def _desc(g, xshape, cout, dtype, algo):
    """xshape = (N, D, H, W, Cin) of the NDHWC forward input."""
    d = ConvDesc()
    n, dd, hh, ww, cin = xshape
    od, oh, ow = g.out_spatial((dd, hh, ww))
    d.x.n, d.x.c, d.x.d, d.x.h, d.x.w = n, cin, dd, hh, ww
    d.y.n, d.y.c, d.y.d, d.y.h, d.y.w = n, cout, od, oh, ow
    for i in range(3):
        d.k[i], d.stride[i], d.pad[i], d.dil[i] = g.k[i], g.stride[i], g.pad[i], g.dil[i]
    d.transposed, d.io_dtype, d.algo = int(g.transposed), _dt(dtype), int(algo)
    return d


def _grad_slot(p):
    from . import parallel
    return parallel.grad_slot(p) if parallel._SLOTS else None


_weights_epoch = [0]
_hook_installed = [False]


def invalidate_packs(optimizer=None, *_a, **_k):
    """REQUIRED after any in-place write to a parameter that does not go through an optimizer step or bump
    ``Tensor._version`` - ``p.data.clamp_()`` weight clipping, a hand-written EMA, ``dist.broadcast(p.data)`` -
    call it without arguments.  (``PACK_DEBUG[0] = True`` re-packs on every cache hit and raises on a stale pack.)

    Drop the cached packed filters of the parameters an optimizer just stepped (called from
    a global optimizer post-step hook: fused / foreach optimizers update parameters in place
    WITHOUT bumping ``Tensor._version``, so the version counter alone cannot be trusted).
    Parameters of other optimizers keep their packs (G's filters stay packed across the
    discriminator steps, trainer.py:296-303).  Without an optimizer argument: drop everything."""
    groups = getattr(optimizer, "param_groups", None)
    if groups is None:
        _weights_epoch[0] += 1
        return
    for g in groups:
        for p in g["params"]:
            if getattr(p, "_b2c_pack", None) is not None:
                p._b2c_pack = None


def watch_optimizers():
    """Register a global optimizer post-step hook that invalidates the packed-filter
    cache.  Idempotent; called by install(), prepare() and GanStep."""
    if not _hook_installed[0]:
        from torch.optim.optimizer import register_optimizer_step_post_hook
        register_optimizer_step_post_hook(invalidate_packs)
        _hook_installed[0] = True


def _packed_filter(w, desc, pas, dtype):
    """bf16/fp32 streaming copy of a torch-layout filter, cached on the tensor
    object and invalidated by its version counter (optimizer steps bump it;
    spectral-norm weights are fresh tensors every forward, utils.py:13)."""
    state = (w.data_ptr(), w._version, _weights_epoch[0])
    cache = getattr(w, "_b2c_pack", None)
    # the packed layout depends on the whole descriptor (one-channel layers switch layout on stride / pad / algo,
    # csrc/thin.cu thin_mode), not only on the pass: the same weight used through F.conv3d with another geometry
    # gets its own pack
    sub = (pas, dtype, int(desc.transposed), tuple(desc.k), tuple(desc.stride), tuple(desc.pad), tuple(desc.dil),
           int(desc.algo))
    if cache is not None and cache[0] == state and sub in cache[1]:
        if PACK_DEBUG[0]:
            _check_pack(w, desc, pas, cache[1][sub])
        return cache[1][sub]
    wf = w.detach()
    if wf.dtype != torch.float32:
        wf = wf.float()
    if not wf.is_contiguous():
        wf = wf.contiguous()
    L = _lib.lib()
    nbytes = L.b2c_filter_packed_bytes(C.byref(desc), pas)
    out = torch.empty(nbytes // (2 if dtype == torch.bfloat16 else 4), dtype=dtype, device=w.device)
    with _dev(w):
        check(L.b2c_filter_pack(C.byref(desc), pas, _ptr(wf), _ptr(out), _stream(w)), "filter_pack")
    try:
        if cache is None or cache[0] != state:
            cache = (state, {})
            w._b2c_pack = cache
        cache[1][sub] = out
    except AttributeError:
        pass
    return out


PACK_DEBUG = [False]     # set PACK_DEBUG[0] = True: every cache hit is re-packed and compared (finds stale packs)


def _check_pack(w, desc, pas, cached):
    L = _lib.lib()
    wf = w.detach().float().contiguous()
    fresh = torch.empty_like(cached)
    with _dev(w):
        check(L.b2c_filter_pack(C.byref(desc), pas, _ptr(wf), _ptr(fresh), _stream(w)), "filter_pack")
    if not torch.equal(fresh, cached):
        raise RuntimeError("b200conv3d: stale packed filter - the parameter was written through `.data` (weight "
                           "clipping, EMA, broadcast) without b200conv3d.functional.invalidate_packs()")


# ------------------------------------------------------------------- conv
class _ConvFn(Function):
    """conv3d / conv_transpose3d with fused bias + activation.

    forward : b2c_conv3d_fwd ; backward: b2c_conv3d_dgrad / b2c_conv3d_wgrad,
    each only when the corresponding input requires grad — the freeze pattern
    of trainer.py:209-210,249-250,255-256 decides which passes run."""

    @staticmethod
    def forward(ctx, x, w, bias, g, act, slope, stats=None):
        """`stats`: None, or a dict {"kind": STAT_CHANNEL | STAT_SAMPLE} that receives "sum" / "sqsum" (fp32) of the
        stored output, accumulated in the GEMM epilogue for the BatchNorm3d / LayerNorm that follows."""
        _check_device(x)
        cd = compute_dtype(x)
        xb = to_ndhwc(x.detach(), cd)
        cout = w.shape[1] if g.transposed else w.shape[0]
        cin_w = w.shape[0] if g.transposed else w.shape[1]
        if cin_w != xb.shape[4]:
            raise RuntimeError("conv3d: expected input with %d channels, got %d (groups != 1 "
                               "is not supported on this path)" % (cin_w, xb.shape[4]))
        desc = _desc(g, tuple(xb.shape), cout, cd, _force_algo[0])
        L = _lib.lib()
        wp = _packed_filter(w, desc, PASS_FWD, cd)
        # a W-folded launch (g.fold > 1) writes [N,D,H,W/F,F*C] = the same bytes as [N,D,H,W,C]: the result is
        # allocated in the UNFOLDED logical shape, so the caller gets a fresh tensor, not a view of one (an in-place
        # op on a view - nn.ReLU(True), `y.add_(x)` - makes autograd rebase it: a clone and two strided copies)
        yl, y = alloc_out(desc.y.n, desc.y.d, desc.y.h, desc.y.w * g.fold, desc.y.c // g.fold, cd, x.device)
        bf = None
        if bias is not None:
            bf = bias.detach()
            bf = bf if bf.dtype == torch.float32 else bf.float()
        fu = Fusion(_ptr(bf), int(act), float(slope), 0, None, None)
        nws = L.b2c_conv3d_workspace(C.byref(desc), PASS_FWD)
        if stats is not None:
            n_st = desc.y.c if stats["kind"] == STAT_CHANNEL else desc.y.n
            st_buf = torch.empty((2, n_st), dtype=torch.float32, device=x.device)
            fu.stat_kind, fu.stat_sum, fu.stat_sqsum = stats["kind"], st_buf[0].data_ptr(), st_buf[1].data_ptr()
            nws = ((nws + 255) & ~255) + L.b2c_conv3d_stats_workspace(C.byref(desc), stats["kind"])
            stats["sum"], stats["sqsum"] = st_buf[0], st_buf[1]
        ws = _ws(nws, x)
        with _dev(x), _timed("fwd", desc, g.fold, g.orig):
            check(L.b2c_conv3d_fwd(C.byref(desc), _ptr(xb), _ptr(wp), _ptr(y), C.byref(fu),
                                   _ptr(ws), ws.numel(), _stream(x)), "conv3d_fwd")
        ctx.g, ctx.act, ctx.slope, ctx.cd = g, int(act), float(slope), cd
        ctx.x_dtype, ctx.has_bias = x.dtype, bias is not None
        ctx.save_for_backward(xb, w, yl if act != ACT_NONE else None)
        return yl

    @staticmethod
    def backward(ctx, gy):
        xb, w, y = ctx.saved_tensors
        if y is not None:
            y = y.permute(0, 2, 3, 4, 1)
        g, cd = ctx.g, ctx.cd
        L = _lib.lib()
        cout = w.shape[1] if g.transposed else w.shape[0]
        desc = _desc(g, tuple(xb.shape), cout, cd, _force_algo[0])
        gyb = to_ndhwc(gy, cd)
        if ctx.act != ACT_NONE:
            gz = torch.empty_like(gyb)
            with _dev(gyb):
                check(L.b2c_act_bwd(_ptr(y), _ptr(gyb), _ptr(gz), gyb.numel(), _dt(cd), ctx.act,
                                    ctx.slope, _stream(gyb)), "act_bwd")
            gyb = gz
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            if _dgrad_merge_applies(g, xb, w, cd):
                dxb = _dgrad_phase_merged(gyb, w, tuple(xb.shape), cd)
            else:
                wp = _packed_filter(w, desc, PASS_DGRAD, cd)
                dxb = torch.empty_like(xb)
                ws = _ws(L.b2c_conv3d_workspace(C.byref(desc), PASS_DGRAD), xb)
                with _dev(xb), _timed("dgrad", desc, g.fold, g.orig):
                    check(L.b2c_conv3d_dgrad(C.byref(desc), _ptr(gyb), _ptr(wp), _ptr(dxb), _ptr(ws),
                                             ws.numel(), _stream(xb)), "conv3d_dgrad")
            dx = from_ndhwc(dxb)
            if ctx.x_dtype == torch.float32 and cd != torch.float32:
                dx = from_ndhwc(cast(dxb, torch.float32))
        need_w = ctx.needs_input_grad[1]
        need_b = ctx.has_bias and ctx.needs_input_grad[2]
        if need_w or need_b:
            # data-parallel runs: the gradient is written straight into the parameter's slot of its all-reduce bucket
            slot = _grad_slot(w) if (need_w and w.dtype == torch.float32) else None
            dw = slot if slot is not None else (torch.empty(w.shape, dtype=torch.float32, device=w.device) if need_w else None)
            db = torch.empty(cout, dtype=torch.float32, device=w.device) if need_b else None
            ws = _ws(L.b2c_conv3d_workspace(C.byref(desc), PASS_WGRAD), xb)
            with _dev(xb), _timed("wgrad", desc, g.fold, g.orig):
                check(L.b2c_conv3d_wgrad(C.byref(desc), _ptr(xb), _ptr(gyb), _ptr(dw), _ptr(db),
                                         _ptr(ws), ws.numel(), _stream(xb)), "conv3d_wgrad")
            if dw is not None and w.dtype != torch.float32:
                dw = dw.to(w.dtype)
        return dx, dw, db, None, None, None, None


def _act_id(act):
    if isinstance(act, int):
        return act
    return ACT_IDS[act]


# True: ask the conv epilogue for statistics wherever a layer wants them; False: never (the norm layers run their own
# pass); "channel" / "sample": only that kind (measurement switch, bench.py --fused-stats)
STATS = [True]
_STAT_IDS = {"channel": STAT_CHANNEL, "sample": STAT_SAMPLE}


def _with_stats(fn_apply, y_args, stats, fold=1):
    """Run a conv Function; attach the statistics its epilogue accumulated to the output tensor (consumed by
    batch_norm_act / layer_norm_act when the tensor reaches them unmodified)."""
    if stats is None or not STATS[0] or (STATS[0] is not True and STATS[0] != stats):
        return fn_apply(*y_args)
    holder = {"kind": _STAT_IDS[stats]}
    y = fn_apply(*y_args, holder)
    if "sum" in holder:
        sm, sq = holder["sum"], holder["sqsum"]
        if fold > 1 and holder["kind"] == STAT_CHANNEL:      # folded channels are (f, c): sum the f copies of c
            sm, sq = sm.view(fold, -1).sum(0), sq.view(fold, -1).sum(0)
        y._b2c_stats = (holder["kind"], sm, sq, y._version)
    return y


def take_stats(x, kind, n):
    """The (sum, sqsum) a producing conv attached to `x`, if they describe its CURRENT contents; else None."""
    st = getattr(x, "_b2c_stats", None)
    if st is None or st[0] != kind or st[3] != x._version or st[1].numel() != n:
        return None
    return st[1], st[2]


def conv3d(x, weight, bias=None, stride=1, padding=0, dilation=1, groups=1, act=None, slope=0.0, stats=None):
    """Drop-in for ``torch.nn.functional.conv3d`` (stylegan.py:168, nn.Conv3d.forward)
    with an optional fused activation epilogue (`act` a name or a b2c activation id) and, with
    ``stats="channel" | "sample"``, the statistics of the output for the BatchNorm3d / LayerNorm that follows."""
    if isinstance(padding, str):
        raise NotImplementedError("b200conv3d.conv3d: string padding %r is not supported" % padding)
    if groups != 1:
        if groups == x.shape[1] and tuple(weight.shape[:2]) == (groups, 1) and bias is None and act is None:
            return _depthwise_fir(x, weight, stride, padding, dilation)
        raise NotImplementedError("b200conv3d.conv3d: groups=%d is not supported (dense convs and the depthwise "
                                  "one-filter FIR of stylegan.py:197 are built; no cuDNN fallback)" % groups)
    g = Geom(False, tuple(weight.shape[2:]), stride, padding, dilation)
    if _s2d_applies(x, weight, g):
        return _conv3d_s2d(x, weight, bias, g, _act_id(act), slope, stats)
    win = _window_factor(x, weight, g)
    if win > 1:
        return _conv3d_windowed(x, weight, bias, g, _act_id(act), slope, win, stats)
    fold = _fold_factor(x, weight, g)
    if fold > 1:
        return _conv3d_folded(x, weight, bias, g, _act_id(act), slope, fold, stats)
    return _with_stats(_ConvFn.apply, (x, weight, bias, g, _act_id(act), slope), stats)


def _depthwise_fir(x, weight, stride, padding, dilation):
    """``F.conv3d(x, f.repeat([C,1,1,1,1]), groups=C)`` of stylegan.py:195-197: every channel is filtered with the SAME
    constant filter, which is what b2c_fir3d computes in one NDHWC pass (its taps are a host array).  One device ->
    host read fetches the taps together with the proof that all channels share them; per-channel filters raise."""
    s, p, d = _triple(stride), _triple(padding), _triple(dilation)
    if len(set(s)) != 1 or d != (1, 1, 1):
        raise NotImplementedError("depthwise conv3d: isotropic stride and dilation 1 only (stylegan.py:197)")
    if weight.requires_grad:
        raise NotImplementedError("depthwise conv3d: the filter is a constant buffer on the reference's path "
                                  "(stylegan.py:462-465); a trainable depthwise filter is not supported")
    k = tuple(weight.shape[2:])
    if k[0] * k[1] * k[2] > 64:
        raise NotImplementedError("depthwise conv3d: at most 64 taps")
    w0 = weight[0, 0].detach().float()
    same = (weight[:, 0].detach().float() == w0).all().float().reshape(1)
    host = torch.cat([w0.reshape(-1), same]).cpu().tolist()
    if host[-1] != 1.0:
        raise NotImplementedError("depthwise conv3d with per-channel filters is not on the reference's path")
    out_sp = tuple((a + 2 * pp - kk) // s[0] + 1 for a, pp, kk in zip(x.shape[2:], p, k))
    if min(out_sp) <= 0:
        raise RuntimeError("conv3d: computed output size is too small")
    return _FirFn.apply(x, host[:-1], k, 1, s[0], p, out_sp)


# ---- space-to-depth for one-channel stride-2 stems with many taps --------------------------------
# FID_ResNet.py:119 `Conv3d(1, 64, kernel_size=7, stride=2, padding=3)`: 343 taps on a one-channel tensor.
# The fused one-channel gather kernel holds <= 64 taps, and the generic lowering writes a 4.6 GB im2col
# matrix (16.8 ms of the 38.8 ms FID forward at batch 25).  Moving the 2x2x2 sub-lattices of the input into
# 8 channels (one pass over a 105 MB tensor) turns the layer into a stride-1 `8 -> Cout` convolution with
# a (k+3)/2-tap "same" kernel (5^3 for k = 7, zero where no original tap lands), which then runs W-folded
# on the tcgen05 path.
S2D = [True]


def _s2d_applies(x, weight, g):
    if not S2D[0] or g.transposed or x.dim() != 5 or weight.shape[1] != 1 or x.shape[1] != 1:
        return False
    k = weight.shape[2]
    if tuple(weight.shape[2:]) != (k, k, k) or k % 2 == 0 or k ** 3 <= 64:
        return False
    if g.stride != (2, 2, 2) or g.dil != (1, 1, 1) or g.pad != (k // 2,) * 3 or weight.shape[0] % 8:
        return False
    return all(v % 2 == 0 for v in x.shape[2:]) and compute_dtype(x) == torch.bfloat16


_s2d_index_cache = {}


def _s2d_index(k, device):
    """idx[p, t] = original tap that connects input voxel 2*(o + t - R) + p to output voxel o (pad = k//2),
    or k (a zero slot); R = the radius of the phase-space kernel."""
    key = (k, str(device))
    if key not in _s2d_index_cache:
        m = k // 2
        r = max(abs((-m) // 2), (m) // 2)
        idx = torch.full((2, 2 * r + 1), k, dtype=torch.long)
        for p in range(2):
            for t in range(2 * r + 1):
                kk = 2 * (t - r) + p + m
                if 0 <= kk < k:
                    idx[p, t] = kk
        _s2d_index_cache[key] = (idx.to(device), r)
    return _s2d_index_cache[key]


def s2d_filter(weight):
    """[Cout, 1, k, k, k] (stride 2, pad k//2) -> [Cout, 8, k', k', k'] (stride 1, pad k'//2); differentiable."""
    cout, _, k = weight.shape[:3]
    idx, r = _s2d_index(k, weight.device)
    kp = 2 * r + 1
    wz = torch.nn.functional.pad(weight[:, 0], (0, 1, 0, 1, 0, 1))                     # zero slot at index k on every axis
    wf = wz[:, idx[:, :, None, None, None, None], idx[None, None, :, :, None, None], idx[None, None, None, None, :, :]]
    return wf.permute(0, 1, 3, 5, 2, 4, 6).reshape(cout, 8, kp, kp, kp), r              # channel = (pd, ph, pw)


def _conv3d_s2d(x, weight, bias, g, act, slope, stats=None):
    n, _, d, h, w = x.shape
    x8 = x.reshape(n, d // 2, 2, h // 2, 2, w // 2, 2).permute(0, 1, 3, 5, 2, 4, 6).reshape(n, d // 2, h // 2, w // 2, 8)
    w8, r = s2d_filter(weight)
    return conv3d(x8.permute(0, 4, 1, 2, 3), w8, bias, 1, r, act=act, slope=slope, stats=stats)


# ---- W-folding of narrow stride-1 convolutions ------------------------------------------------
# The tcgen05 conv kernel fetches one TMA row per (voxel, tap) and is bound by the NUMBER of
# L2 -> SM row requests, not by their bytes (a 3^3 conv re-fetches every input row 27 times):
# measured, a 32-channel row (64 B) costs the same as a 128-channel row (256 B), so 32 -> 32
# layers at 128^3 (utils.py:57-60,105-109 hidden convs, stylegan.py conv0 / conv1 at the top
# resolution) run at 145 TFLOP/s where 128 -> 128 layers reach 1190.  Folding F adjacent voxels
# along W into the channel axis is FREE in NDHWC storage ([N,D,H,W,C] == [N,D,H,W/F,F*C]) and
# turns the layer into a (F*Cin) -> (F*Cout) convolution over W/F with 3 taps along W whose
# filter has a known block-sparsity pattern: F x fewer row requests for (at most) F x the
# tensor-core work, which the kernel has to spare.  Weight gradients come back through the
# (index_select) expansion: every original tap is the sum of the blocks it was copied to.
FOLD = [True]
FOLD_K1 = [True]
FOLD_1CH = [False]     # measured: the folded 8 <-> 8*C GEMMs leave the tensor-core path (K = 8) - 278 -> 408 ms/step


def _fold_factor(x, weight, g):
    if not FOLD[0] or g.transposed or x.dim() != 5:
        return 1
    cout, cin, kd, kh, kw = weight.shape
    if g.stride != (1, 1, 1) or g.dil != (1, 1, 1) or kw % 2 == 0 or g.pad[2] != kw // 2:
        return 1
    # experiment knob (off): one-channel 1x1x1 convs (fromrgb / torgb, stylegan.py:542-546,758) folded by 8 into
    # 8 <-> 8*C block-diagonal GEMMs.  The forward of torgb gains (1.07 -> 0.22 ms) but the K = 8 weight / data
    # gradients fall to the CUDA-core kernels; the one-channel lowering of csrc/thin.cu stays the default.
    if FOLD_1CH[0] and (kd, kh, kw) == (1, 1, 1) and min(cin, cout) == 1 and max(cin, cout) % 8 == 0 \
            and max(cin, cout) <= 32 and x.shape[4] % 8 == 0 and compute_dtype(x) == torch.bfloat16:
        return 8
    # 1x1x1 convs (utils.py:55,60,101,111) are pure row traffic: fold up to 64 channels (block-diagonal filter)
    cmax = 64 if (kd, kh, kw) == (1, 1, 1) else 32
    thin_in = cin <= 16 and cout <= 64                      # 16- / 32-byte input rows: fold by 4 whatever the output
    if cin % 8 or cout % 8 or (max(cin, cout) > cmax and not thin_in) or compute_dtype(x) != torch.bfloat16 \
            or not FOLD_K1[0] and kw == 1:
        return 1
    w = x.shape[4]
    for f in ((4, 2) if (max(cin, cout) <= cmax // 2 or thin_in) else (2,)):
        if w % f == 0 and w // f >= 2 and kw // 2 <= f:
            return f
    return 1


_fold_index_cache = {}


def _fold_index(f, kw, device):
    """idx[po, pi, t] = original tap along W that connects input voxel F*(j+t-1)+pi to output voxel
    F*j+po (pad = kw//2), or kw (a zero slot) if there is none."""
    key = (f, kw, str(device))
    if key not in _fold_index_cache:
        pad = kw // 2
        nt, t0 = (1, 0) if kw == 1 else (3, 1)          # folded taps along W and the centre tap
        idx = torch.full((f, f, nt), kw, dtype=torch.long)
        for po in range(f):
            for pi in range(f):
                for t in range(nt):
                    k = f * (t - t0) + pi - po + pad
                    if 0 <= k < kw:
                        idx[po, pi, t] = k
        _fold_index_cache[key] = idx.reshape(-1).to(device)
    return _fold_index_cache[key]


def fold_filter(weight, f):
    """[Cout, Cin, kd, kh, kw] -> [F*Cout, F*Cin, kd, kh, 3] (block-sparse; differentiable); a 1x1x1 filter
    becomes the block-diagonal [F*Cout, F*Cin, 1, 1, 1]."""
    cout, cin, kd, kh, kw = weight.shape
    wz = torch.cat([weight, weight.new_zeros(cout, cin, kd, kh, 1)], dim=4)
    nt = 1 if kw == 1 else 3
    wf = wz.index_select(4, _fold_index(f, kw, weight.device)).reshape(cout, cin, kd, kh, f, f, nt)
    return wf.permute(4, 0, 5, 1, 2, 3, 6).reshape(f * cout, f * cin, kd, kh, nt)


def _fold_w(x, f):
    """logical [N,C,D,H,W] -> logical [N,F*C,D,H,W/F] over the same NDHWC storage (a view)."""
    n, c, d, h, w = x.shape
    p = x.permute(0, 2, 3, 4, 1)
    if not p.is_contiguous():
        p = p.contiguous()
    return p.reshape(n, d, h, w // f, f * c).permute(0, 4, 1, 2, 3)


def _unfold_w(y, f):
    n, fc, d, h, wf = y.shape
    return y.permute(0, 2, 3, 4, 1).reshape(n, d, h, wf * f, fc // f).permute(0, 4, 1, 2, 3)


def _conv3d_folded(x, weight, bias, g, act, slope, f, stats=None):
    w2 = fold_filter(weight, f)
    b2 = None if bias is None else bias.repeat(f)
    g2 = Geom(False, tuple(w2.shape[2:]), 1, (g.pad[0], g.pad[1], 1 if w2.shape[4] == 3 else 0), 1)
    g2.fold = f
    return _with_stats(_ConvFn.apply, (_fold_w(x, f), w2, b2, g2, act, slope), stats, f)   # unfolded by _ConvFn


# ---- W-window lowering of narrow 3-tap convolutions (second generation of the folding above) -------------------
# Folding F voxels into the channel axis keeps THREE taps along W, of which the outer two use one of their F
# voxel blocks: 1/2 (F = 2) or 1/4 (F = 4) of the issued MACs are useful, and with F*Cout = 64 half of the 128-row
# UMMA tile is padding.  Gathering, per group of F output voxels, the F + 2 input voxels it needs
# (b2c_window_w: [rows, W, C] -> [rows, W/F, (F+2)*C], one extra HBM pass over 1.5x the input) leaves ONE tap along W:
# a (F+2)*Cin -> F*Cout convolution with a block-Toeplitz filter, 3 / (F + 2) of its MACs useful, and F = 128 / Cout
# fills the tile: 32 -> 32 at 128^3 becomes 192 -> 128 with 9 taps (half of the MACs useful, 6.75 row requests per
# output voxel instead of 13.5), 64 -> 64 at 64^3 becomes 256 -> 128 (three quarters useful).  The weight gradient
# comes back through the index_select that builds the filter, the data gradient through the adjoint gather.
WINDOW = [True]


def _window_factor(x, weight, g):
    if not WINDOW[0] or g.transposed or x.dim() != 5:
        return 1
    cout, cin, kd, kh, kw = weight.shape
    if g.stride != (1, 1, 1) or g.dil != (1, 1, 1) or kw != 3 or g.pad[2] != 1 or cout not in (32, 64):
        return 1
    f = 128 // cout
    if cin % 8 or cin < 16 or (f + 2) * cin > 512 or x.shape[4] % f or x.shape[4] // f < 2:
        return 1
    if compute_dtype(x) != torch.bfloat16 or not (x.is_cuda or _lib.is_host()):
        return 1
    return f


_win_index_cache = {}


def _win_index(f, device):
    """idx[vo, vi] = tap kw = vi - vo that connects window voxel vi to output voxel vo of the group, or 3 (a zero slot)."""
    key = (f, str(device))
    if key not in _win_index_cache:
        idx = torch.full((f, f + 2), 3, dtype=torch.long)
        for vo in range(f):
            for k in range(3):
                idx[vo, vo + k] = k
        _win_index_cache[key] = idx.reshape(-1).to(device)
    return _win_index_cache[key]


def window_filter(weight, f):
    """[Cout, Cin, kd, kh, 3] -> block-Toeplitz [F*Cout, (F+2)*Cin, kd, kh, 1] (differentiable)."""
    cout, cin, kd, kh, _ = weight.shape
    wz = torch.cat([weight, weight.new_zeros(cout, cin, kd, kh, 1)], dim=4)
    wf = wz.index_select(4, _win_index(f, weight.device)).reshape(cout, cin, kd, kh, f, f + 2)
    return wf.permute(4, 0, 5, 1, 2, 3).reshape(f * cout, (f + 2) * cin, kd, kh, 1)


class _WindowWFn(Function):
    """x [N,C,D,H,W] -> [N,(F+2)*C,D,H,W/F]: the F+2 input voxels of every group of F output voxels (b2c_window_w)."""

    @staticmethod
    def forward(ctx, x, f):
        _check_device(x)
        cd = compute_dtype(x)
        xb = to_ndhwc(x.detach(), cd)
        n, d, h, w, c = xb.shape
        yl, y = alloc_out(n, d, h, w // f, (f + 2) * c, cd, x.device)
        with _dev(x):
            check(_lib.lib().b2c_window_w(_ptr(xb), _ptr(y), n * d * h, w, c, f, 1, _dt(cd), 0, _stream(x)), "window_w")
        ctx.meta = (n, d, h, w, c, f, cd, x.dtype)
        return yl

    @staticmethod
    def backward(ctx, gy):
        n, d, h, w, c, f, cd, x_dtype = ctx.meta
        g = to_ndhwc(gy, cd)
        dxl, dx = alloc_out(n, d, h, w, c, cd, gy.device)
        with _dev(gy):
            check(_lib.lib().b2c_window_w(_ptr(g), _ptr(dx), n * d * h, w, c, f, 1, _dt(cd), 1, _stream(gy)), "window_w")
        return (dxl if x_dtype == cd else dxl.to(x_dtype)), None


def _conv3d_windowed(x, weight, bias, g, act, slope, f, stats=None):
    cout, cin = weight.shape[0], weight.shape[1]
    w2 = window_filter(weight, f)
    b2 = None if bias is None else bias.repeat(f)
    g2 = Geom(False, tuple(w2.shape[2:]), 1, (g.pad[0], g.pad[1], 0), 1)
    g2.fold, g2.orig = f, (cin, cout)
    xw = _WindowWFn.apply(x, f)
    return _with_stats(_ConvFn.apply, (xw, w2, b2, g2, act, slope), stats, f)        # unfolded by _ConvFn (same storage)


def conv_transpose3d(x, weight, bias=None, stride=1, padding=0, output_padding=0, groups=1,
                     dilation=1, act=None, slope=0.0, stats=None):
    """Drop-in for ``torch.nn.functional.conv_transpose3d`` (stylegan.py:170,
    nn.ConvTranspose3d.forward).  Computed as a stride-decomposed sub-pixel
    convolution inside the library, never by zero insertion."""
    if groups != 1:
        raise NotImplementedError("b200conv3d.conv_transpose3d: groups=%d is not supported" % groups)
    g = Geom(True, tuple(weight.shape[2:]), stride, padding, dilation, output_padding)
    return _with_stats(_ConvFn.apply, (x, weight, bias, g, _act_id(act), slope), stats)


# ---- phase-merged stride-2 transposed conv for narrow outputs -------------------------------------
# The up-sampling layers of stylegan.py:262-282 are `conv_transpose3d(k = 3, stride 2)` with 32 / 64 output
# channels.  In the library's scatter form that is 8 sub-pixel problems of 1 .. 8 taps with N = Cout <= 64
# columns each: one 128-byte TMA row per 64 x 32 MACs, 77 TFLOP/s on 64 -> 32 at 64^3 -> 129^3.  The 8 phases
# read the SAME input voxels (offsets 0 / -1 per axis), so they are merged into one stride-1 convolution
# `Cin -> 8*Cout` with a 2x2x2 kernel (zero where a phase has no tap: 27 of 64 slots are live) on the
# (in + 1)^3 lattice, followed by ONE depth-to-space pass.  2.4x the MACs, at the tensor pipe's wide-N rate.
# The result is returned on the even-sized grid 2*(in + 1): its last plane / row / column is exactly zero
# (odd phase of the last lattice point), so the FIR that follows takes one voxel less of trailing padding
# instead of a cropped (strided) tensor.
PHASE_MERGE = [True]

_pm_index_cache = {}


def _pm_index(device):
    """idx[r, u] = tap kd*9 + kh*3 + kw of the 3^3 transposed-conv filter that connects input voxel j + u - 1 to
    output voxel 2*j + r (per axis: kappa = 2*(1 - u) + r), or 27 (a zero slot)."""
    key = str(device)
    if key not in _pm_index_cache:
        idx = torch.full((8, 8), 27, dtype=torch.long)
        for r in range(8):
            for u in range(8):
                ks = [2 * (1 - ((u >> s) & 1)) + ((r >> s) & 1) for s in (2, 1, 0)]
                if all(0 <= k < 3 for k in ks):
                    idx[r, u] = (ks[0] * 3 + ks[1]) * 3 + ks[2]
        _pm_index_cache[key] = idx.reshape(-1).to(device)
    return _pm_index_cache[key]


def phase_merge_applies(x, weight):
    """weight: conv_transpose3d layout [Cin, Cout, 3, 3, 3]; stride 2, padding 0."""
    if not PHASE_MERGE[0] or x.dim() != 5 or tuple(weight.shape[2:]) != (3, 3, 3):
        return False
    cin, cout = weight.shape[0], weight.shape[1]
    return cin % 8 == 0 and cout % 8 == 0 and cin >= 16 and 8 * cout <= 512 and x.is_cuda \
        and compute_dtype(x) == torch.bfloat16


def phase_merged_filter(weight):
    """[Cin, Cout, 3,3,3] (transposed-conv layout) -> [8*Cout, Cin, 2,2,2] (conv layout), phase-major channels."""
    cin, cout = weight.shape[0], weight.shape[1]
    w = weight.permute(1, 0, 2, 3, 4).reshape(cout, cin, 27)
    w = torch.cat([w, w.new_zeros(cout, cin, 1)], dim=2)
    w = w.index_select(2, _pm_index(weight.device)).reshape(cout, cin, 8, 8)
    return w.permute(2, 0, 1, 3).reshape(8 * cout, cin, 2, 2, 2)


def _d2s_raw(src, shape_out, c, dt, inverse):
    """src NDHWC-contiguous; z side [n,jd,jh,jw,8c], y side [n,od,oh,ow,c]; shape_out = the destination's."""
    dst = torch.empty(shape_out, dtype=src.dtype, device=src.device)
    zs, ys = (shape_out, src.shape) if inverse else (src.shape, shape_out)
    with _dev(src):
        check(_lib.lib().b2c_depth_to_space2(_ptr(src), _ptr(dst), zs[0], zs[1], zs[2], zs[3], c, ys[1], ys[2], ys[3],
                                             dt, 1 if inverse else 0, _stream(src)), "depth_to_space2")
    return dst


class _D2SFn(Function):
    """z [N, 8*C, J, J, J] (phase-major channels) -> y [N, C, od, oh, ow], y[2j + r] = z[j, r], cropped to `out`."""

    @staticmethod
    def forward(ctx, z, out):
        _check_device(z)
        cd = _keep_dtype(z)
        zb = to_ndhwc(z.detach(), cd)
        n, jd, jh, jw, c8 = zb.shape
        ctx.meta = (tuple(zb.shape), cd, z.dtype)
        yb = _d2s_raw(zb, (n, out[0], out[1], out[2], c8 // 8), c8 // 8, _dt(cd), False)
        return from_ndhwc(yb)

    @staticmethod
    def backward(ctx, gy):
        zshape, cd, z_dtype = ctx.meta
        gz = from_ndhwc(_d2s_raw(to_ndhwc(gy, cd), zshape, zshape[4] // 8, _dt(cd), True))
        return (gz if z_dtype == cd else gz.to(z_dtype)), None


def conv_transpose3d_s2(x, weight, even=False):
    """`conv_transpose3d(x, weight, stride=2, padding=0)` for a 3^3 filter as ONE stride-1 `Cin -> 8*Cout` conv with
    a 2x2x2 kernel on the (in + 1)^3 lattice + depth-to-space (b2c_depth_to_space2).  `even`: return the even grid
    2*(in + 1) per axis, whose last plane / row / column is exactly zero, instead of the 2*in + 1 result."""
    d, h, w = x.shape[2:]
    z = conv3d(x, phase_merged_filter(weight), None, 1, 1)                  # [N, 8*Cout, d+1, h+1, w+1]
    return _D2SFn.apply(z, tuple(2 * v + (2 if even else 1) for v in (d, h, w)))


def _dgrad_phase_merged(gyb, w, dshape, cd):
    """dgrad of `conv3d(k = 3, stride 2, padding 0)` with a narrow input (stylegan.py:257-260 down-sampling convs:
    32@129^3 -> 64@64^3 ran its sub-pixel scatter form at 77 TFLOP/s): the same merged conv + depth-to-space,
    written straight into the odd-sized dx.  gyb NDHWC [N, o, o, o, Cout]; w [Cout, Cin, 3,3,3]; returns NDHWC dx."""
    with torch.no_grad():
        z = conv3d(from_ndhwc(gyb), phase_merged_filter(w.detach().float()), None, 1, 1)
        zb = to_ndhwc(z, cd)
        return _d2s_raw(zb, dshape, dshape[4], _dt(cd), False)


def _dgrad_merge_applies(g, xb, w, cd):
    if not PHASE_MERGE[0] or g.transposed or tuple(w.shape[2:]) != (3, 3, 3) or cd != torch.bfloat16:
        return False
    if g.stride != (2, 2, 2) or g.pad != (0, 0, 0) or g.dil != (1, 1, 1):
        return False
    cout, cin = w.shape[0], w.shape[1]
    # every input voxel must lie on the even grid of the (out + 1)^3 lattice: in <= 2 * (out + 1)
    return cin % 8 == 0 and cout % 8 == 0 and cout >= 16 and 8 * cin <= 512 and xb.is_cuda


# -------------------------------------------------------------- batch norm
def _f32(t):
    if t is None:
        return None
    t = t.detach()
    return t if t.dtype == torch.float32 and t.is_contiguous() else t.float().contiguous()


class _BNActFn(Function):
    """BatchNorm3d (+ residual add) (+ activation) in training or eval mode.
    dcgan.py:48-64, utils.py:62-86, biggan.py:40-41, FID_ResNet.py:71-100."""

    @staticmethod
    def forward(ctx, x, gamma, beta, residual, rmean, rvar, training, momentum, eps, act, slope, pre=None):
        """`pre`: (sum, sqsum) per channel delivered by the producing conv's epilogue, or None."""
        _check_device(x)
        cd = compute_dtype(x)
        xb = to_ndhwc(x.detach(), cd)
        rb = to_ndhwc(residual.detach(), cd) if residual is not None else None
        n, d, h, w, c = xb.shape
        rows = n * d * h * w
        L = _lib.lib()
        dev = x.device
        stats = torch.empty((6, c), dtype=torch.float32, device=dev)
        s_sum, s_sq, mean, invstd, scale, shift = stats.unbind(0)
        gf, bf = _f32(gamma), _f32(beta)
        with _dev(x):
            st = _stream(x)
            if training:
                if pre is not None:
                    s_sum, s_sq = pre[0], pre[1]
                else:
                    ws = _ws(L.b2c_stats_workspace(rows, c), x)
                    check(L.b2c_channel_stats(_ptr(xb), rows, c, _dt(cd), _ptr(s_sum), _ptr(s_sq),
                                              _ptr(ws), ws.numel(), st), "channel_stats")
                check(L.b2c_bn_finalize(_ptr(s_sum), _ptr(s_sq), rows, c, float(eps),
                                        float(momentum), _ptr(gf), _ptr(bf), _ptr(mean),
                                        _ptr(invstd), _ptr(scale), _ptr(shift),
                                        _ptr(rmean), _ptr(rvar), st), "bn_finalize")
            else:
                # eval mode: [C]-sized glue on the running statistics
                mean.copy_(rmean)
                invstd.copy_(torch.rsqrt(rvar.float() + eps))
                scale.copy_(invstd if gf is None else gf * invstd)
                shift.copy_((0 if bf is None else bf) - mean * scale)
            yl, y = alloc_out(n, d, h, w, c, cd, dev)
            check(L.b2c_bn_act_fwd(_ptr(xb), _ptr(y), _ptr(rb), rows, c, _dt(cd), _ptr(scale),
                                   _ptr(shift), int(act), float(slope), st), "bn_act_fwd")
        ctx.save_for_backward(xb, rb, stats)
        ctx.training, ctx.act, ctx.slope, ctx.cd = bool(training), int(act), float(slope), cd
        ctx.x_dtype = x.dtype
        ctx.res_dtype = residual.dtype if residual is not None else None
        ctx.affine = (gamma is not None, beta is not None)
        return yl

    @staticmethod
    def backward(ctx, gy):
        xb, rb, stats = ctx.saved_tensors
        _, _, mean, invstd, scale, shift = stats.unbind(0)
        cd = ctx.cd
        n, d, h, w, c = xb.shape
        rows = n * d * h * w
        L = _lib.lib()
        gyb = to_ndhwc(gy, cd)
        s = torch.zeros((2, c), dtype=torch.float32, device=xb.device)
        s1, s2 = s.unbind(0)
        need_x = ctx.needs_input_grad[0]
        need_res = rb is not None and ctx.needs_input_grad[3]
        need_aff = ctx.needs_input_grad[1] or ctx.needs_input_grad[2]
        dx = dres = None
        with _dev(xb):
            st = _stream(xb)
            if ctx.training or need_aff:
                ws = _ws(L.b2c_stats_workspace(rows, c), xb)
                check(L.b2c_bn_act_bwd_reduce(_ptr(xb), _ptr(gyb), _ptr(rb), rows, c, _dt(cd),
                                              _ptr(mean), _ptr(invstd), _ptr(scale), _ptr(shift),
                                              ctx.act, ctx.slope, _ptr(s1), _ptr(s2), _ptr(ws),
                                              ws.numel(), st), "bn_act_bwd_reduce")
            if need_x or need_res:
                dxb = torch.empty_like(xb)
                drb = torch.empty_like(xb) if need_res else None
                # eval mode: statistics are constants -> no mean-subtraction terms
                a1, a2 = (s1, s2) if ctx.training else (torch.zeros_like(s1), torch.zeros_like(s2))
                check(L.b2c_bn_act_bwd_apply(_ptr(xb), _ptr(gyb), _ptr(rb), _ptr(dxb), _ptr(drb),
                                             rows, c, _dt(cd), _ptr(mean), _ptr(invstd),
                                             _ptr(scale), _ptr(shift), _ptr(a1), _ptr(a2),
                                             ctx.act, ctx.slope, st), "bn_act_bwd_apply")
                if need_x:
                    dx = from_ndhwc(cast(dxb, torch.float32) if ctx.x_dtype == torch.float32 else dxb)
                if need_res:
                    dres = from_ndhwc(cast(drb, torch.float32) if ctx.res_dtype == torch.float32 else drb)
        dgamma = s2.clone() if (ctx.affine[0] and ctx.needs_input_grad[1]) else None
        dbeta = s1.clone() if (ctx.affine[1] and ctx.needs_input_grad[2]) else None
        return dx, dgamma, dbeta, dres, None, None, None, None, None, None, None, None


def batch_norm_act(x, gamma, beta, running_mean, running_var, training, momentum, eps,
                   act=None, slope=0.0, residual=None):
    use_batch = training or running_mean is None
    pre = take_stats(x, STAT_CHANNEL, x.shape[1]) if (use_batch and x.dim() == 5) else None
    return _BNActFn.apply(x, gamma, beta, residual, running_mean, running_var, use_batch,
                          0.0 if momentum is None else momentum, eps, _act_id(act), slope, pre)


# -------------------------------------------------------------- layer norm
def _perm_affine(p, c, s, to_ndhwc_order, out=None):
    """[C,D,H,W] fp32 <-> [D,H,W,C] fp32 through b2c_transpose.  The parameter -> NDHWC copy is
    cached on the parameter like a packed filter (same invalidation rules).  `out`: destination (flat view)."""
    state = (p.data_ptr(), p._version, _weights_epoch[0])
    cache = getattr(p, "_b2c_pack", None) if to_ndhwc_order else None
    if cache is not None and cache[0] == state and ("ln", c, s) in cache[1]:
        return cache[1][("ln", c, s)]
    pf = _f32(p).reshape(-1)
    out = torch.empty_like(pf) if out is None else out.view(-1)
    rows, cols = (c, s) if to_ndhwc_order else (s, c)
    with _dev(pf):
        check(_lib.lib().b2c_transpose(_ptr(pf), _ptr(out), 1, rows, cols, F32, F32, _stream(pf)),
              "transpose")
    if to_ndhwc_order:
        try:
            if cache is None or cache[0] != state:
                cache = (state, {})
                p._b2c_pack = cache
            cache[1][("ln", c, s)] = out
        except AttributeError:
            pass
    return out


class _LNActFn(Function):
    """nn.LayerNorm([C,D,H,W]) (+ LeakyReLU) of the DCGAN discriminator,
    dcgan.py:162-179: per-sample statistics over C*D*H*W and a full-shape
    elementwise affine."""

    @staticmethod
    def forward(ctx, x, gamma, beta, eps, act, slope, pre=None):
        """`pre`: (sum, sqsum) per sample delivered by the producing conv's epilogue, or None."""
        _check_device(x)
        cd = compute_dtype(x)
        xb = to_ndhwc(x.detach(), cd)
        n, d, h, w, c = xb.shape
        Lh = d * h * w * c
        L = _lib.lib()
        gp = _perm_affine(gamma, c, d * h * w, True) if gamma is not None else None
        bp = _perm_affine(beta, c, d * h * w, True) if beta is not None else None
        mr = torch.empty((2, n), dtype=torch.float32, device=x.device)
        yl, y = alloc_out(n, d, h, w, c, cd, x.device)
        if pre is not None:
            with _dev(x):
                check(L.b2c_ln_act_fwd_stats(_ptr(xb), _ptr(y), n, Lh, _dt(cd), _ptr(gp), _ptr(bp), float(eps),
                                             int(act), float(slope), _ptr(pre[0]), _ptr(pre[1]), _ptr(mr[0]),
                                             _ptr(mr[1]), _stream(x)), "ln_act_fwd_stats")
        else:
            ws = _ws(L.b2c_ln_workspace(n, Lh), x)
            with _dev(x):
                check(L.b2c_ln_act_fwd(_ptr(xb), _ptr(y), n, Lh, _dt(cd), _ptr(gp), _ptr(bp),
                                       float(eps), int(act), float(slope), _ptr(mr[0]), _ptr(mr[1]),
                                       _ptr(ws), ws.numel(), _stream(x)), "ln_act_fwd")
        ctx.save_for_backward(xb, gp, bp, mr)
        ctx.act, ctx.slope, ctx.cd, ctx.x_dtype = int(act), float(slope), cd, x.dtype
        ctx.pshape = tuple(gamma.shape) if gamma is not None else None
        ctx.affine = (gamma, beta)
        return yl

    @staticmethod
    def backward(ctx, gy):
        xb, gp, bp, mr = ctx.saved_tensors
        cd = ctx.cd
        n, d, h, w, c = xb.shape
        Lh = d * h * w * c
        L = _lib.lib()
        gyb = to_ndhwc(gy, cd)
        need_x = ctx.needs_input_grad[0]
        need_g = gp is not None and ctx.needs_input_grad[1]
        need_b = bp is not None and ctx.needs_input_grad[2]
        dxb = torch.empty_like(xb) if need_x else None
        dg = torch.empty(Lh, dtype=torch.float32, device=xb.device) if need_g else None
        db = torch.empty(Lh, dtype=torch.float32, device=xb.device) if need_b else None
        ws = _ws(L.b2c_ln_workspace(n, Lh), xb)
        with _dev(xb):
            check(L.b2c_ln_act_bwd(_ptr(xb), _ptr(gyb), _ptr(dxb), n, Lh, _dt(cd), _ptr(gp),
                                   _ptr(bp), _ptr(mr[0]), _ptr(mr[1]), ctx.act, ctx.slope,
                                   _ptr(dg), _ptr(db), _ptr(ws), ws.numel(), _stream(xb)),
                  "ln_act_bwd")
        dx = None
        if need_x:
            dx = from_ndhwc(cast(dxb, torch.float32) if ctx.x_dtype == torch.float32 else dxb)
        def unperm(grad, param):          # NDHWC order -> the parameter's [C,D,H,W], into its all-reduce slot if any
            slot = _grad_slot(param) if param.dtype == torch.float32 else None
            return _perm_affine(grad, c, d * h * w, False, out=slot).reshape(ctx.pshape) if slot is None else \
                (_perm_affine(grad, c, d * h * w, False, out=slot), slot)[1]
        dgamma = unperm(dg, ctx.affine[0]) if need_g else None
        dbeta = unperm(db, ctx.affine[1]) if need_b else None
        return dx, dgamma, dbeta, None, None, None, None


def layer_norm_act(x, gamma, beta, eps, act=None, slope=0.0):
    pre = take_stats(x, STAT_SAMPLE, x.shape[0]) if x.dim() == 5 else None
    return _LNActFn.apply(x, gamma, beta, eps, _act_id(act), slope, pre)


# --------------------------------------------------------------- pointwise
def _dense_flat(x):
    """(flat dense view or None) for tensors whose storage is dense in some
    permutation — elementwise kernels do not care about the order."""
    if x.is_contiguous():
        return x.view(-1)
    if x.dim() == 5:
        p = x.permute(0, 2, 3, 4, 1)
        if p.is_contiguous():
            return p.reshape(-1)
    return None


class _ActFn(Function):
    """ReLU / LeakyReLU / Tanh (dcgan.py:49,68,163; utils.py:68; biggan.py:63)."""

    @staticmethod
    def forward(ctx, x, act, slope, inplace):
        _check_device(x)
        L = _lib.lib()
        flat = _dense_flat(x) if x.dtype in (torch.bfloat16, torch.float32) else None
        if inplace and flat is not None:
            ctx.mark_dirty(x)
            out, oflat = x, flat
        else:
            if flat is None:
                xs = x.contiguous()
                xs = xs if xs.dtype in (torch.bfloat16, torch.float32) else xs.float()
                flat = xs.view(-1)
                out = torch.empty_like(xs)
                oflat = out.view(-1)
            else:
                out = torch.empty_like(x)      # preserves (dense) strides
                oflat = _dense_flat(out)
        with _dev(x):
            check(L.b2c_act_fwd(_ptr(flat), _ptr(oflat), flat.numel(), _dt(flat.dtype), int(act),
                                float(slope), _stream(x)), "act_fwd")
        if inplace and out is not x:
            x.copy_(out)
            ctx.mark_dirty(x)
            out = x
        ctx.act, ctx.slope = int(act), float(slope)
        ctx.save_for_backward(out)
        return out

    @staticmethod
    def backward(ctx, gy):
        (y,) = ctx.saved_tensors
        L = _lib.lib()
        yf = _dense_flat(y)
        if yf is None or y.dtype not in (torch.bfloat16, torch.float32):
            y = y.contiguous()
            y = y if y.dtype in (torch.bfloat16, torch.float32) else y.float()
            yf = y.view(-1)
        # bring gy to y's physical order and dtype
        g = gy if gy.dtype == y.dtype else gy.to(y.dtype)
        if g.stride() != y.stride() or _dense_flat(g) is None:
            g2 = torch.empty_like(y)
            g2.copy_(g)
            g = g2
        gf = _dense_flat(g)
        dx = torch.empty_like(y)
        with _dev(y):
            check(L.b2c_act_bwd(_ptr(yf), _ptr(gf), _ptr(_dense_flat(dx)), yf.numel(),
                                _dt(yf.dtype), ctx.act, ctx.slope, _stream(y)), "act_bwd")
        return dx, None, None, None


def activation(x, act, slope=0.0, inplace=False):
    return _ActFn.apply(x, _act_id(act), slope, inplace)


# --------------------------------------------------------------- x2 resampling (NDHWC)
def _keep_dtype(x):
    """bf16 / fp32 tensors keep their dtype; fp16 (the reference's autocast dtype, trainer.py:213) maps to bf16 like
    compute_dtype does, so the resampling passes between two convs stay 2-byte; anything else is plumbing -> fp32."""
    if x.dtype in (torch.bfloat16, torch.float32):
        return x.dtype
    return torch.bfloat16 if x.dtype == torch.float16 else torch.float32


class _Up2Fn(Function):
    """`x[:, :channels]` + F.interpolate(scale_factor=2) (nearest) of utils.py:76-80 in one
    NDHWC pass; backward sums 2x2x2 blocks and zero-fills the dropped channels."""

    @staticmethod
    def forward(ctx, x, channels):
        _check_device(x)
        cd = _keep_dtype(x)
        xb = to_ndhwc(x.detach(), cd)
        n, d, h, w, cx = xb.shape
        c = cx if channels is None else int(channels)
        if not 0 < c <= cx:
            raise ValueError("channels=%r out of range for a %d-channel tensor" % (channels, cx))
        yl, y = alloc_out(n, 2 * d, 2 * h, 2 * w, c, cd, x.device)
        with _dev(x):
            check(_lib.lib().b2c_upsample2x_fwd(_ptr(xb), _ptr(y), n, d, h, w, cx, 0, c, _dt(cd), _stream(x)),
                  "upsample2x_fwd")
        ctx.geo, ctx.cd, ctx.x_dtype = (n, d, h, w, cx, c), cd, x.dtype
        return yl

    @staticmethod
    def backward(ctx, gy):
        n, d, h, w, cx, c = ctx.geo
        gyb = to_ndhwc(gy, ctx.cd)
        dxl, dx = alloc_out(n, d, h, w, cx, ctx.cd, gy.device)
        with _dev(gy):
            check(_lib.lib().b2c_upsample2x_bwd(_ptr(gyb), _ptr(dx), n, d, h, w, cx, 0, c, _dt(ctx.cd),
                                                _stream(gy)), "upsample2x_bwd")
        return (dxl if ctx.x_dtype == ctx.cd else dxl.to(ctx.x_dtype)), None


def upsample_nearest2x(x, channels=None):
    """Drop-in for ``F.interpolate(x[:, :channels], scale_factor=2)`` (nearest) on 5-D tensors."""
    return _Up2Fn.apply(x, channels)


class _AvgPool2Fn(Function):
    """nn.AvgPool3d(2) (utils.py:113-116,128-129; biggan.py:86,92) on NDHWC storage."""

    @staticmethod
    def forward(ctx, x):
        _check_device(x)
        cd = _keep_dtype(x)
        xb = to_ndhwc(x.detach(), cd)
        n, d, h, w, c = xb.shape
        if min(d, h, w) < 2:
            raise ValueError("AvgPool3d(2) needs every spatial extent >= 2, got %s" % ((d, h, w),))
        yl, y = alloc_out(n, d // 2, h // 2, w // 2, c, cd, x.device)
        with _dev(x):
            check(_lib.lib().b2c_avgpool2_fwd(_ptr(xb), _ptr(y), n, d, h, w, c, _dt(cd), _stream(x)), "avgpool2_fwd")
        ctx.geo, ctx.cd, ctx.x_dtype = (n, d, h, w, c), cd, x.dtype
        return yl

    @staticmethod
    def backward(ctx, gy):
        n, d, h, w, c = ctx.geo
        gyb = to_ndhwc(gy, ctx.cd)
        dxl, dx = alloc_out(n, d, h, w, c, ctx.cd, gy.device)
        with _dev(gy):
            check(_lib.lib().b2c_avgpool2_bwd(_ptr(gyb), _ptr(dx), n, d, h, w, c, _dt(ctx.cd), _stream(gy)),
                  "avgpool2_bwd")
        return dxl if ctx.x_dtype == ctx.cd else dxl.to(ctx.x_dtype)


def avg_pool2(x):
    """Drop-in for ``nn.AvgPool3d(2)(x)``."""
    return _AvgPool2Fn.apply(x)


# --------------------------------------------------------------- depthwise FIR resampling
class _FirFn(Function):
    """One NDHWC pass for pad / zero-insert / depthwise FIR / decimate (stylegan.py:172-200)."""

    @staticmethod
    def forward(ctx, x, taps, k, up, down, pad_lo, out_sp):
        _check_device(x)
        cd = _keep_dtype(x)
        xb = to_ndhwc(x.detach(), cd)
        n, d, h, w, c = xb.shape
        desc = _lib.FirDesc()
        desc.n, desc.c, desc.up, desc.down, desc.dtype = n, c, up, down, _dt(cd)
        for i, (a, o) in enumerate(zip((d, h, w), out_sp)):
            desc.in_[i], desc.out[i], desc.k[i], desc.pad[i] = a, o, k[i], pad_lo[i]
        tarr = (C.c_float * len(taps))(*taps)
        yl, y = alloc_out(n, out_sp[0], out_sp[1], out_sp[2], c, cd, x.device)
        with _dev(x):
            check(_lib.lib().b2c_fir3d_fwd(C.byref(desc), tarr, _ptr(xb), _ptr(y), _stream(x)), "fir3d_fwd")
        ctx.desc, ctx.tarr, ctx.cd, ctx.x_dtype, ctx.in_shape = desc, tarr, cd, x.dtype, (n, d, h, w, c)
        return yl

    @staticmethod
    def backward(ctx, gy):
        n, d, h, w, c = ctx.in_shape
        gyb = to_ndhwc(gy, ctx.cd)
        dxl, dx = alloc_out(n, d, h, w, c, ctx.cd, gy.device)
        with _dev(gy):
            check(_lib.lib().b2c_fir3d_bwd(C.byref(ctx.desc), ctx.tarr, _ptr(gyb), _ptr(dx), _stream(gy)), "fir3d_bwd")
        return (dxl if ctx.x_dtype == ctx.cd else dxl.to(ctx.x_dtype)), None, None, None, None, None, None


def upfirdn3d(x, f, up=1, down=1, padding=0, flip_filter=False, gain=1):
    """Drop-in for ``stylegan._upfirdn3d_ref`` (stylegan.py:172-200) for the isotropic arguments the
    reference uses: `f` a 3-D filter (or None), integer `up` / `down`, `padding` an int, 3 or 6 ints
    in the reference's F.pad order.  The filter carries no gradient (a constant buffer there)."""
    if isinstance(up, (list, tuple)) or isinstance(down, (list, tuple)):
        if len(set(up if isinstance(up, (list, tuple)) else [up])) != 1 or \
                len(set(down if isinstance(down, (list, tuple)) else [down])) != 1:
            raise NotImplementedError("anisotropic up / down factors are not on the reference's path")
        up = up[0] if isinstance(up, (list, tuple)) else up
        down = down[0] if isinstance(down, (list, tuple)) else down
    p = padding
    if isinstance(p, int):
        p = [p] * 6
    elif len(p) == 3:
        p = [p[0], p[0], p[1], p[1], p[2], p[2]]
    p = [int(v) for v in p]
    if len(p) != 6:
        raise ValueError("padding must be an int, 3 ints or 6 ints")
    if min(p) < 0 and (len(set(p[0::2])) != 1 or len(set(p[1::2])) != 1):
        # stylegan.py:186-187 pads (dim4, dim3, dim2) but crops (dim3, dim2, dim4): only defined alike
        raise NotImplementedError("anisotropic negative padding (crop) is not on the reference's path")
    if f is None:
        taps, k = [1.0], (1, 1, 1)
    else:
        if f.dim() != 3:
            raise NotImplementedError("upfirdn3d takes the reference's 3-D (4x4x4) filters")
        ff = f.detach().float().cpu() * float(gain)
        if not flip_filter:
            ff = ff.flip([0, 1, 2])
        taps, k = ff.reshape(-1).tolist(), tuple(ff.shape)
    # F.pad order: (p[0], p[1]) -> dim 4, (p[2], p[3]) -> dim 3, (p[4], p[5]) -> dim 2
    lo = (p[4], p[2], p[0])
    hi = (p[5], p[3], p[1])
    out_sp = []
    for a, l, hh, kk in zip(x.shape[2:], lo, hi, k):
        full = a * up + l + hh - kk + 1
        if full <= 0:
            raise ValueError("upfirdn3d: empty output")
        out_sp.append(-(-full // down))
    return _FirFn.apply(x, taps, k, int(up), int(down), lo, tuple(out_sp))


# --------------------------------------------------------------- style modulation glue (NDHWC)
def mod_supported(c, dtype):
    """b2c_modulate_* / b2c_mod_epilogue_* take whole 16-byte channel vectors, at most 256 of them."""
    vec = 8 if dtype == torch.bfloat16 else 4
    return c % vec == 0 and c // vec <= 256


def _f32c(t):
    return None if t is None else t.detach().to(torch.float32).contiguous()


class _ModulateFn(Function):
    """xs = x * s[n, c]  (stylegan.py:427); backward: dx = g * s, ds[n,c] = sum_v g * x, one pass."""

    @staticmethod
    def forward(ctx, x, s):
        _check_device(x)
        cd = compute_dtype(x)
        xb = to_ndhwc(x.detach(), cd)
        n, d, h, w, c = xb.shape
        sf = _f32c(s).reshape(n, c)
        yl, y = alloc_out(n, d, h, w, c, cd, x.device)
        with _dev(x):
            check(_lib.lib().b2c_modulate_fwd(_ptr(xb), _ptr(sf), _ptr(y), n, d * h * w, c, _dt(cd), _stream(x)),
                  "modulate_fwd")
        ctx.save_for_backward(xb, sf)
        ctx.x_dtype, ctx.s_dtype, ctx.s_shape, ctx.cd = x.dtype, s.dtype, s.shape, cd
        return yl

    @staticmethod
    def backward(ctx, gy):
        xb, sf = ctx.saved_tensors
        n, d, h, w, c = xb.shape
        g = to_ndhwc(gy, ctx.cd)
        dxl, dx = alloc_out(n, d, h, w, c, ctx.cd, gy.device)
        ds = torch.empty(n, c, dtype=torch.float32, device=gy.device)
        L = _lib.lib()
        ws = _ws(L.b2c_mod_workspace(n, d * h * w, c), gy)
        with _dev(gy):
            check(L.b2c_modulate_bwd(_ptr(xb), _ptr(g), _ptr(sf), _ptr(dx), _ptr(ds), n, d * h * w, c, _dt(ctx.cd),
                                     _ptr(ws), ws.numel(), _stream(gy)), "modulate_bwd")
        return (dxl if ctx.x_dtype == ctx.cd else dxl.to(ctx.x_dtype)), ds.reshape(ctx.s_shape).to(ctx.s_dtype)


def modulate(x, s):
    """x [N,C,D,H,W] * s [N,C] broadcast over the voxels; one NDHWC pass each way."""
    return _ModulateFn.apply(x, s)


class _ModEpilogueFn(Function):
    """out = act(y * d[n,c] + strength * field[n|0, v] + bias[c])  (stylegan.py:429-432, 522-524, 531);
    backward: dy = g * d, dd[n,c] = sum_v g * y, dbias[c] = sum g, dstrength = sum g * field with
    g = gout * act'(out) - one pass over (y, out, gout) instead of a dozen ATen passes."""

    @staticmethod
    def forward(ctx, y, d, field, strength, bias, act, slope):
        _check_device(y)
        cd = compute_dtype(y)
        yb = to_ndhwc(y.detach(), cd)
        n, dd_, h, w, c = yb.shape
        v = dd_ * h * w
        df, bf = _f32c(d), _f32c(bias)
        ff = sf = None
        fn = 0
        if field is not None:
            ff = _f32c(field).reshape(-1, v)
            fn = ff.shape[0]
            if fn not in (1, n):
                raise ValueError("noise field batch must be 1 or N")
            sf = _f32c(strength).reshape(1)
        outl, out = alloc_out(n, dd_, h, w, c, cd, y.device)
        with _dev(y):
            check(_lib.lib().b2c_mod_epilogue_fwd(_ptr(yb), _ptr(df), _ptr(ff), fn, _ptr(sf), _ptr(bf), _ptr(out), n, v,
                                                  c, _dt(cd), int(act), float(slope), _stream(y)), "mod_epilogue_fwd")
        need_y = d is not None and ctx.needs_input_grad[1]
        ctx.save_for_backward(yb if need_y else None, out, df, ff)
        ctx.meta = (cd, y.dtype, int(act), float(slope), fn,
                    None if d is None else (d.shape, d.dtype),
                    None if strength is None else (strength.shape, strength.dtype),
                    None if bias is None else bias.dtype)
        return outl

    @staticmethod
    def backward(ctx, gout):
        yb, out, df, ff = ctx.saved_tensors
        cd, y_dtype, act, slope, fn, dmeta, smeta, bdtype = ctx.meta
        n, d_, h, w, c = out.shape
        v = d_ * h * w
        g = to_ndhwc(gout, cd)
        dyl, dy = alloc_out(n, d_, h, w, c, cd, gout.device)
        dev = gout.device
        want_d = yb is not None
        want_b = bdtype is not None and ctx.needs_input_grad[4]
        want_s = ff is not None and smeta is not None and ctx.needs_input_grad[3]
        dd = torch.empty(n, c, dtype=torch.float32, device=dev) if want_d else None
        db = torch.empty(c, dtype=torch.float32, device=dev) if want_b else None
        dst = torch.empty(1, dtype=torch.float32, device=dev) if want_s else None
        L = _lib.lib()
        ws = _ws(L.b2c_mod_workspace(n, v, c), gout)
        with _dev(gout):
            check(L.b2c_mod_epilogue_bwd(_ptr(yb), _ptr(out), _ptr(g), _ptr(df), _ptr(ff), fn, _ptr(dy), _ptr(dd),
                                         _ptr(db), _ptr(dst), n, v, c, _dt(cd), act, slope, _ptr(ws), ws.numel(),
                                         _stream(gout)), "mod_epilogue_bwd")
        return ((dyl if y_dtype == cd else dyl.to(y_dtype)),
                None if dd is None else dd.reshape(dmeta[0]).to(dmeta[1]),
                None,
                None if dst is None else dst.reshape(smeta[0]).to(smeta[1]),
                None if db is None else db.to(bdtype),
                None, None)


def mod_epilogue(y, d=None, field=None, strength=None, bias=None, act=None, slope=0.0):
    """y [N,C,D,H,W]; d [N,C] or None; field [N|1, 1, D,H,W] (or anything of N*V / V elements) with its
    scalar `strength` (a tensor), or None; bias [C] or None."""
    return _ModEpilogueFn.apply(y, d, field, strength, bias, _act_id(act), slope)


# --------------------------------------------------------------- AdaIN (+ trilinear x2), v1-like generator
class _AdaINFn(Function):
    """y = (up ? tri2x : id)((x - mean) / sqrt(var_unbiased + eps)) * s_std[n,c] + s_mean[n,c]
    (stylegan.py:931-959 followed by F.interpolate(scale_factor=2, mode='trilinear'), :1081).
    Forward: one statistics pass + one apply pass that reads the low-resolution tensor and writes the
    (up-sampled) output.  Backward: one reduce pass (adjoint interpolation + per-(n,c) sums) and one
    apply pass; the [N,C]-sized algebra between them is done in fp64 on the device by ATen."""

    @staticmethod
    def forward(ctx, x, s_mean, s_std, eps, up):
        _check_device(x)
        cd = compute_dtype(x)
        xb = to_ndhwc(x.detach(), cd)
        n, d, h, w, c = xb.shape
        v = d * h * w
        L = _lib.lib()
        s1 = torch.empty(n, c, dtype=torch.float32, device=x.device)
        s2 = torch.empty_like(s1)
        ws = _ws(L.b2c_mod_workspace(n, v * (8 if up else 1), c), x)
        with _dev(x):
            check(L.b2c_instance_stats(_ptr(xb), _ptr(s1), _ptr(s2), n, v, c, _dt(cd), _ptr(ws), ws.numel(),
                                       _stream(x)), "instance_stats")
        mean = s1.double() / v
        var = (s2.double() - v * mean * mean).clamp_min(0.0) / max(v - 1, 1)          # torch.var: unbiased
        sigma = (var + eps).sqrt()
        a64 = s_std.detach().double().reshape(n, c) / sigma
        b64 = s_mean.detach().double().reshape(n, c) - mean * a64
        a, b = a64.float().contiguous(), b64.float().contiguous()
        f = 2 if up else 1
        yl, y = alloc_out(n, d * f, h * f, w * f, c, cd, x.device)
        with _dev(x):
            check(L.b2c_adain_fwd(_ptr(xb), _ptr(a), _ptr(b), _ptr(y), n, d, h, w, c, _dt(cd), int(up), _stream(x)),
                  "adain_fwd")
        ctx.save_for_backward(xb, mean, sigma, s_std.detach().double().reshape(n, c))
        ctx.meta = (cd, x.dtype, bool(up), s_mean.shape, s_mean.dtype, s_std.shape, s_std.dtype)
        return yl

    @staticmethod
    def backward(ctx, gy):
        xb, mean, sigma, S = ctx.saved_tensors
        cd, x_dtype, up, m_shape, m_dtype, s_shape, s_dtype = ctx.meta
        n, d, h, w, c = xb.shape
        v = d * h * w
        g = to_ndhwc(gy, cd)
        L = _lib.lib()
        dxl, dx = alloc_out(n, d, h, w, c, cd, gy.device)
        r1 = torch.empty(n, c, dtype=torch.float32, device=gy.device)
        r2 = torch.empty_like(r1)
        ws = _ws(L.b2c_mod_workspace(n, v, c), gy)
        with _dev(gy):
            check(L.b2c_adain_bwd_reduce(_ptr(xb), _ptr(g), _ptr(dx) if up else None, _ptr(r1), _ptr(r2), n, d, h, w, c,
                                         _dt(cd), int(up), _ptr(ws), ws.numel(), _stream(gy)), "adain_bwd_reduce")
        r1d, r2d = r1.double(), r2.double()
        gx = (r2d - mean * r1d) / sigma                   # sum_i t_i * xhat_i  = d L / d s_std
        # dx_i = (S/sigma) [t_i - r1/V - xhat_i * gx / (V-1)],  xhat_i = (x_i - mean) / sigma
        ct = S / sigma
        cx = -ct * gx / (sigma * max(v - 1, 1))
        c0 = -ct * r1d / v - cx * mean
        t = dx if up else g
        ctf, cxf, c0f = ct.float().contiguous(), cx.float().contiguous(), c0.float().contiguous()   # keep alive
        with _dev(gy):
            check(L.b2c_adain_bwd_apply(_ptr(xb), _ptr(t), _ptr(ctf), _ptr(cxf), _ptr(c0f), _ptr(dx), n, v, c,
                                        _dt(cd), _stream(gy)), "adain_bwd_apply")
        return ((dxl if x_dtype == cd else dxl.to(x_dtype)), r1d.reshape(m_shape).to(m_dtype),
                gx.reshape(s_shape).to(s_dtype), None, None)


def adain(x, s_mean, s_std, eps=1e-5, up=False):
    """Instance-normalise x [N,C,D,H,W] per (n, c) (unbiased variance, as torch.var), scale by s_std [N,C] and
    shift by s_mean [N,C]; with up=True the trilinear x2 interpolation that follows is folded into the pass."""
    return _AdaINFn.apply(x, s_mean, s_std, float(eps), bool(up))


# --------------------------------------------------------------- MaxPool3d on NDHWC
class _MaxPoolFn(Function):
    """nn.MaxPool3d(k, stride, pad) (FID_ResNet.py:129) in one NDHWC pass (ATen converts to NCDHW and back)."""

    @staticmethod
    def forward(ctx, x, k, stride, pad):
        _check_device(x)
        cd = _keep_dtype(x)
        xb = to_ndhwc(x.detach(), cd)
        n, d, h, w, c = xb.shape
        od, oh, ow = ((v + 2 * pad - k) // stride + 1 for v in (d, h, w))
        yl, y = alloc_out(n, od, oh, ow, c, cd, x.device)
        with _dev(x):
            check(_lib.lib().b2c_maxpool3d_fwd(_ptr(xb), _ptr(y), n, d, h, w, c, k, stride, pad, _dt(cd), _stream(x)),
                  "maxpool3d_fwd")
        ctx.save_for_backward(xb)
        ctx.meta = (k, stride, pad, cd, x.dtype)
        return yl

    @staticmethod
    def backward(ctx, gy):
        (xb,) = ctx.saved_tensors
        k, stride, pad, cd, x_dtype = ctx.meta
        n, d, h, w, c = xb.shape
        g = to_ndhwc(gy, cd)
        dxl, dx = alloc_out(n, d, h, w, c, cd, gy.device)
        with _dev(gy):
            check(_lib.lib().b2c_maxpool3d_bwd(_ptr(xb), _ptr(g), _ptr(dx), n, d, h, w, c, k, stride, pad, _dt(cd),
                                               _stream(gy)), "maxpool3d_bwd")
        return (dxl if x_dtype == cd else dxl.to(x_dtype)), None, None, None


def max_pool3d(x, kernel_size, stride=None, padding=0):
    """Drop-in for F.max_pool3d with one isotropic int per argument (what the reference uses)."""
    k, s, p = _triple(kernel_size), _triple(stride if stride is not None else kernel_size), _triple(padding)
    if len(set(k)) != 1 or len(set(s)) != 1 or len(set(p)) != 1:
        raise NotImplementedError("anisotropic MaxPool3d is not on the reference's path")
    return _MaxPoolFn.apply(x, k[0], s[0], p[0])
