"""ctypes binding of the C ABI in include/b2c.h.

The product library is ``_lib/libb2c.so`` next to this file (built in-tree by
``__graft_entry__.build()`` from ``3d-gan_b200/csrc``).  There is NO fallback:
if the library is missing, importing an op raises.  Tests (and only tests) may
point the binding at the CPU oracle with :func:`use_library` to exercise the
host logic in a container without a GPU.
"""
import ctypes as C
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
PRODUCT_LIB = os.path.join(_HERE, "_lib", "libb2c.so")

# constants mirrored from include/b2c.h
BF16, F32 = 0, 2
ACT_NONE, ACT_RELU, ACT_LRELU, ACT_TANH = 0, 1, 2, 3
PASS_FWD, PASS_DGRAD, PASS_WGRAD = 0, 1, 2
ALGO_AUTO, ALGO_DIRECT, ALGO_TC = 0, 1, 2
ACT_IDS = {None: ACT_NONE, "none": ACT_NONE, "relu": ACT_RELU,
           "lrelu": ACT_LRELU, "tanh": ACT_TANH}


class Shape(C.Structure):
    _fields_ = [(k, C.c_int32) for k in ("n", "c", "d", "h", "w")]


class ConvDesc(C.Structure):
    _fields_ = [("x", Shape), ("y", Shape),
                ("k", C.c_int32 * 3), ("stride", C.c_int32 * 3),
                ("pad", C.c_int32 * 3), ("dil", C.c_int32 * 3),
                ("transposed", C.c_int32), ("io_dtype", C.c_int32),
                ("algo", C.c_int32)]


class FirDesc(C.Structure):
    """b2c_fir_desc (include/b2c.h)."""
    _fields_ = [("n", C.c_int64), ("c", C.c_int32), ("in_", C.c_int32 * 3), ("out", C.c_int32 * 3),
                ("k", C.c_int32 * 3), ("pad", C.c_int32 * 3), ("up", C.c_int32), ("down", C.c_int32),
                ("dtype", C.c_int32)]


STAT_NONE, STAT_CHANNEL, STAT_SAMPLE = 0, 1, 2


class Fusion(C.Structure):
    """b2c_fusion (include/b2c.h); the three trailing fields default to "no statistics"."""
    _fields_ = [("bias", C.c_void_p), ("act", C.c_int32), ("slope", C.c_float), ("stat_kind", C.c_int32),
                ("stat_sum", C.c_void_p), ("stat_sqsum", C.c_void_p)]


_vp, _i64, _i32, _f32, _sz = C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_size_t
_D = C.POINTER(ConvDesc)
_SIGNATURES = {
    "b2c_version": (C.c_int, []),
    "b2c_strerror": (C.c_char_p, [_i32]),
    "b2c_last_algo": (C.c_int, []),
    "b2c_launch_count": (C.c_uint64, []),
    "b2c_filter_packed_bytes": (_sz, [_D, _i32]),
    "b2c_filter_pack": (_i32, [_D, _i32, _vp, _vp, _vp]),
    "b2c_conv3d_workspace": (_sz, [_D, _i32]),
    "b2c_conv3d_stats_workspace": (_sz, [_D, _i32]),
    "b2c_conv3d_fwd": (_i32, [_D, _vp, _vp, _vp, C.POINTER(Fusion), _vp, _sz, _vp]),
    "b2c_conv3d_dgrad": (_i32, [_D, _vp, _vp, _vp, _vp, _sz, _vp]),
    "b2c_conv3d_wgrad": (_i32, [_D, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "b2c_transpose": (_i32, [_vp, _vp, _i64, _i64, _i64, _i32, _i32, _vp]),
    "b2c_cast": (_i32, [_vp, _vp, _i64, _i32, _i32, _vp]),
    "b2c_stats_workspace": (_sz, [_i64, _i64]),
    "b2c_channel_stats": (_i32, [_vp, _i64, _i64, _i32, _vp, _vp, _vp, _sz, _vp]),
    "b2c_channel_stats_static": (_i32, [_vp, _i64, _i64, _i32, _vp, _vp, _vp]),
    "b2c_bn_finalize": (_i32, [_vp, _vp, _i64, _i64, _f32, _f32, _vp, _vp, _vp, _vp,
                               _vp, _vp, _vp, _vp, _vp]),
    "b2c_bn_act_fwd": (_i32, [_vp, _vp, _vp, _i64, _i64, _i32, _vp, _vp, _i32, _f32, _vp]),
    "b2c_bn_act_bwd_reduce": (_i32, [_vp, _vp, _vp, _i64, _i64, _i32, _vp, _vp, _vp, _vp,
                                     _i32, _f32, _vp, _vp, _vp, _sz, _vp]),
    "b2c_bn_act_bwd_apply": (_i32, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _i32, _vp, _vp,
                                    _vp, _vp, _vp, _vp, _i32, _f32, _vp]),
    "b2c_ln_workspace": (_sz, [_i64, _i64]),
    "b2c_ln_act_fwd": (_i32, [_vp, _vp, _i64, _i64, _i32, _vp, _vp, _f32, _i32, _f32,
                              _vp, _vp, _vp, _sz, _vp]),
    "b2c_ln_act_fwd_stats": (_i32, [_vp, _vp, _i64, _i64, _i32, _vp, _vp, _f32, _i32, _f32, _vp, _vp, _vp, _vp, _vp]),
    "b2c_ln_act_bwd": (_i32, [_vp, _vp, _vp, _i64, _i64, _i32, _vp, _vp, _vp, _vp,
                              _i32, _f32, _vp, _vp, _vp, _sz, _vp]),
    "b2c_act_fwd": (_i32, [_vp, _vp, _i64, _i32, _i32, _f32, _vp]),
    "b2c_act_bwd": (_i32, [_vp, _vp, _vp, _i64, _i32, _i32, _f32, _vp]),
    "b2c_upsample2x_fwd": (_i32, [_vp, _vp, _i64, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _vp]),
    "b2c_upsample2x_bwd": (_i32, [_vp, _vp, _i64, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _vp]),
    "b2c_avgpool2_fwd": (_i32, [_vp, _vp, _i64, _i32, _i32, _i32, _i32, _i32, _vp]),
    "b2c_avgpool2_bwd": (_i32, [_vp, _vp, _i64, _i32, _i32, _i32, _i32, _i32, _vp]),
    "b2c_fir3d_fwd": (_i32, [C.POINTER(FirDesc), C.POINTER(C.c_float), _vp, _vp, _vp]),
    "b2c_fir3d_bwd": (_i32, [C.POINTER(FirDesc), C.POINTER(C.c_float), _vp, _vp, _vp]),
    "b2c_mod_workspace": (_sz, [_i64, _i64, _i64]),
    "b2c_modulate_fwd": (_i32, [_vp, _vp, _vp, _i64, _i64, _i64, _i32, _vp]),
    "b2c_modulate_bwd": (_i32, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _i32, _vp, _sz, _vp]),
    "b2c_mod_epilogue_fwd": (_i32, [_vp, _vp, _vp, _i64, _vp, _vp, _vp, _i64, _i64, _i64, _i32, _i32, _f32, _vp]),
    "b2c_mod_epilogue_bwd": (_i32, [_vp, _vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _i32, _i32,
                                    _f32, _vp, _sz, _vp]),
    "b2c_instance_stats": (_i32, [_vp, _vp, _vp, _i64, _i64, _i64, _i32, _vp, _sz, _vp]),
    "b2c_adain_fwd": (_i32, [_vp, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _i64, _i32, _i32, _vp]),
    "b2c_adain_bwd_reduce": (_i32, [_vp, _vp, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _i64, _i32, _i32, _vp, _sz, _vp]),
    "b2c_adain_bwd_apply": (_i32, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _i32, _vp]),
    "b2c_maxpool3d_fwd": (_i32, [_vp, _vp, _i64, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _vp]),
    "b2c_maxpool3d_bwd": (_i32, [_vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _vp]),
    "b2c_window_w": (_i32, [_vp, _vp, _i64, _i32, _i32, _i32, _i32, _i32, _i32, _vp]),
    "b2c_peer_alloc": (_i32, [_sz, C.POINTER(_vp), C.c_char_p]),
    "b2c_peer_open": (_i32, [C.c_char_p, C.POINTER(_vp)]),
    "b2c_peer_close": (_i32, [_vp]),
    "b2c_peer_free": (_i32, [_vp]),
    "b2c_peer_allreduce_avg": (_i32, [C.POINTER(_vp), C.POINTER(_vp), _i32, _i32, _i64, C.c_uint32, _i32, _vp]),
    "b2c_peer_error": (C.c_uint, []),
    "b2c_depth_to_space2": (_i32, [_vp, _vp, _i64, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _vp]),
}
EXPORTS = tuple(_SIGNATURES)

_lock = threading.Lock()
_state = {"lib": None, "host": False, "path": None}


def _bind(path):
    lib = C.CDLL(path)
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if a symbol is missing
        fn.restype, fn.argtypes = res, args
    return lib


def use_library(path, host=False):
    """Bind to an explicit shared library.  ``host=True`` declares that it
    takes HOST pointers (the CPU oracle) — test infrastructure only."""
    with _lock:
        _state.update(lib=_bind(path), host=bool(host), path=path)


def lib():
    if _state["lib"] is None:
        with _lock:
            if _state["lib"] is None:
                if not os.path.exists(PRODUCT_LIB):
                    raise RuntimeError(
                        "b200conv3d: %s is missing — run __graft_entry__.build() "
                        "(there is no CPU or cuDNN fallback)" % PRODUCT_LIB)
                _state.update(lib=_bind(PRODUCT_LIB), host=False, path=PRODUCT_LIB)
    return _state["lib"]


def is_host():
    lib()
    return _state["host"]


def library_path():
    lib()
    return _state["path"]


class B2CError(RuntimeError):
    pass


def check(rc, what):
    if rc != 0:
        msg = lib().b2c_strerror(int(rc)).decode()
        exc = ValueError if rc == -1 else (NotImplementedError if rc == -2 else B2CError)
        raise exc("b200conv3d.%s failed: %s (code %d)" % (what, msg, rc))
