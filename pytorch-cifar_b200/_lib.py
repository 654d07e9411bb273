"""ctypes binding of libb2c.so — the C-ABI declared in include/b2c.h.

The shared library is built in-tree by `csrc/Makefile` (see `build()`); there is no CPU or library
fallback: if the library is missing, importing this module raises.
"""
import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
# B2C_LIB: measurement aid (tools/), an alternative build of the same library; the product path is the in-tree file
LIB_PATH = os.environ.get("B2C_LIB") or os.path.join(_HERE, "libb2c.so")
CSRC = os.path.join(_HERE, "csrc")

B2C_ERR_UNSUPPORTED = -10001
B2C_ERR_DRIVER = -10002
B2C_ERR_ARG = -10003
B2C_ERR_TOO_LARGE = -10004


class B2cError(RuntimeError):
    def __init__(self, code, where):
        self.code = code
        msg = _lib.b2c_error_string(code).decode() if _lib is not None else str(code)
        super().__init__(f"{where}: {msg} (code {code})")


class ConvDesc(ctypes.Structure):
    """Mirror of B2cConvDesc (include/b2c.h)."""
    _fields_ = [(n, ctypes.c_int) for n in
                ("n", "h", "w", "cin", "cout", "ksize", "stride", "cin_valid", "cout_valid", "kflat_cin", "block_n", "max_ctas",
                 "flags", "x_ld", "y_ld")]


class BnStats(ctypes.Structure):
    """Mirror of B2cBnStats (include/b2c.h)."""
    _fields_ = [(n, ctypes.c_void_p) for n in
                ("sum", "sumsq", "gamma", "beta", "conv_bias", "running_mean", "running_var", "scale", "shift",
                 "save_mean", "save_rstd")] + [(n, ctypes.c_float) for n in ("count", "eps", "momentum")]


class RepackEntry(ctypes.Structure):
    """Mirror of B2cRepackEntry (include/b2c.h)."""
    _fields_ = [("src_off", ctypes.c_longlong), ("wf_off", ctypes.c_longlong), ("wd_off", ctypes.c_longlong),
                ("cout", ctypes.c_int), ("cin", ctypes.c_int), ("taps", ctypes.c_int), ("cin_pad", ctypes.c_int),
                ("kflat", ctypes.c_int), ("pad_", ctypes.c_int)]


def build(verbose=False):
    """Compile libb2c.so for sm_100a with nvcc (cross-compiles without a GPU)."""
    r = subprocess.run(["make", "-C", CSRC, "-j8"], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout[-4000:])
        print(r.stderr[-4000:])
    if r.returncode != 0:
        raise RuntimeError("building libb2c.so failed")
    return LIB_PATH


_P = ctypes.c_void_p
_I = ctypes.c_int
_F = ctypes.c_float
_LL = ctypes.c_longlong

# name -> argtypes; every function returns int except b2c_error_string
SIGNATURES = {
    "b2c_version": [],
    "b2c_conv_fprop": [ctypes.POINTER(ConvDesc), _P, _P, _P, _P, _P, _P, ctypes.POINTER(_I), _P],
    "b2c_conv1x1_cat_fprop": [ctypes.POINTER(ConvDesc), _P, _P, _P, _P, _P, _P, ctypes.POINTER(_I), _P],
    "b2c_conv1x1_catn_fprop": [ctypes.POINTER(ConvDesc), ctypes.POINTER(ctypes.c_void_p), _I, _P, _P, _P, _P,
                               ctypes.POINTER(_I), _P],
    "b2c_conv_dgrad": [ctypes.POINTER(ConvDesc), _P, _P, _P, _P, _P],
    "b2c_conv_dgrad_bn": [ctypes.POINTER(ConvDesc), _P, _P, _P, _P, _P, _P, _P, _P, ctypes.POINTER(ctypes.c_int), _P],
    "b2c_conv_dgrad_with_shortcut": [ctypes.POINTER(ConvDesc), _P, _P, _P, _P, _P, _P],
    "b2c_conv_wgrad": [ctypes.POINTER(ConvDesc), _P, _P, _P, _P, _LL, _P],
    "b2c_dwconv_stat_rows": [_LL, _I],
    "b2c_dwconv_fprop": [_P, _P, _P, _I, _I, _I, _I, _I, _I, _P, _P, ctypes.POINTER(_I), _P],
    "b2c_dwconv_input_transform_ok": [_I, _I, _I, _I],
    "b2c_dwconv_fprop_bnrelu": [_P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _P, _P, ctypes.POINTER(_I), _I, _P],
    "b2c_dwconv_wgrad_bnrelu": [_P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _P],
    "b2c_dwconv_dgrad": [_P, _P, _P, _I, _I, _I, _I, _I, _I, _P],
    "b2c_dwconv_wgrad": [_P, _P, _P, _I, _I, _I, _I, _I, _I, _P],
    "b2c_bn_finalize": [_P, _P, _I, _I, _F, _P, _P, _P, _F, _F, _P, _P, _P, _P, _P, _P, _P],
    "b2c_bn_eval_scale_shift": [_I, _P, _P, _P, _P, _P, _F, _P, _P, _P],
    "b2c_bn_act_apply": [_P, _P, _P, _P, _P, _P, _P, _LL, _I, _I, _P, _P],
    "b2c_bn_act_apply_stats": [_P, ctypes.POINTER(BnStats), _P, ctypes.POINTER(BnStats), _P, _LL, _I, _I, _P, _P],
    "b2c_bn_bwd_reduce_acc": [_P, _P, _P, _P, _LL, _I, _P, _P, _P, _P, _P],
    "b2c_bn_bwd_apply_stats": [_P, _P, _P, _F, _P, _P, _P, _P, _P, _P, _P, _LL, _I, _P, _P, _P],
    "b2c_dwconv_fprop_acc": [_P, _P, _P, _I, _I, _I, _I, _I, _I, _P, _P, _P],
    "b2c_bn_bwd_partial_rows": [_LL, _I],
    "b2c_bn_bwd_reduce": [_P, _P, _P, _P, _LL, _I, _P, _P, _P, _P, _P],
    "b2c_bn_bwd_finalize": [_P, _I, _I, _F, _P, _P, _P, _P, _P, _P, _P, _P, _P],
    "b2c_bn_bwd_apply": [_P, _P, _P, _P, _P, _P, _P, _LL, _I, _P, _P, _P],
    "b2c_maxpool2x2_fwd": [_P, _P, _I, _I, _I, _I, _P],
    "b2c_maxpool2x2_bwd": [_P, _P, _P, _P, _I, _I, _I, _I, _P],
    "b2c_bn_bwd_stats_noalloc": [_P, _P, _P, _P, _LL, _I, _F, _P, _P, _P, _P, _P, _P, _P, _P, _P],
    "b2c_head_fwd": [_P, _I, _I, _I, _P, _P, _I, _P, _P, _P],
    "b2c_ce_loss": [_P, _P, _I, _I, _P, _P, _P, _P],
    "b2c_maxpool3x3_fwd": [_P, _P, _P, _I, _I, _I, _I, _I, _P],
    "b2c_maxpool3x3_bwd": [_P, _P, _P, _P, _I, _I, _I, _I, _I, _P],
    "b2c_copy_channels": [_P, _I, _I, _P, _I, _I, _LL, _I, _P],
    "b2c_conv1x1_catr_fprop": [ctypes.POINTER(ConvDesc), ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(_I), _I, _P, _P,
                               _P, _P, ctypes.POINTER(_I), _P],
    "b2c_se_fwd": [_P, _I, _I, _I, _I, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P],
    "b2c_se_bwd": [_P, _P, _I, _I, _I, _I, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P],
    "b2c_head_fwd_ld": [_P, _I, _I, _I, _I, _P, _P, _I, _P, _P, _P],
    "b2c_head_bwd_ld": [_P, _P, _P, _I, _I, _I, _I, _I, _P, _P, _P, _P],
    "b2c_head_bwd": [_P, _P, _P, _I, _I, _I, _I, _P, _P, _P, _P],
    "b2c_augment_nchw": [_P, _P, _P, _P, _I, ctypes.POINTER(_F), ctypes.POINTER(_F), _P],
    "b2c_augment_im2col": [_P, _P, _P, _P, _I, ctypes.POINTER(_F), ctypes.POINTER(_F), _P],
    "b2c_stem_im2col": [_P, _P, _I, _P],
    "b2c_augment_nhwc64": [_P, _P, _P, _P, _I, ctypes.POINTER(_F), ctypes.POINTER(_F), _P],
    "b2c_stem_nhwc64": [_P, _P, _I, _P],
    "b2c_grouped_wgrad_extract": [_P, _P, _I, _I, _I, _I, _P],
    "b2c_sgd_step": [_P, _P, _P, _LL, _P, _P],
    "b2c_repack_weights": [_P, _P, _P, _P, _I, _P],
}

_lib = None
recording = None   # when a list: every C-ABI call is appended as (name, args) before it is issued


class _LibProxy:
    """Attribute access returns the ctypes function wrapped so calls can be recorded and replayed
    (bench.py uses this to count kernel launches and to time kernel families in isolation)."""

    def __init__(self, cdll):
        self._cdll = cdll
        self._cache = {}

    def __getattr__(self, name):
        fn = getattr(self._cdll, name)
        if name in ("b2c_error_string", "b2c_version", "b2c_bn_bwd_partial_rows", "b2c_dwconv_stat_rows",
                    "b2c_dwconv_input_transform_ok"):
            return fn

        def call(*args):
            if recording is not None:
                recording.append((name, args))
            return fn(*args)
        self.__dict__[name] = call
        return call

    def raw(self, name):
        return getattr(self._cdll, name)


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            f"or `make -C {CSRC}`. There is no fallback path.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, args in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = _I
    lib.b2c_error_string.argtypes = [_I]
    lib.b2c_error_string.restype = ctypes.c_char_p
    _lib = _LibProxy(lib)
    return _lib


def check(code, where):
    if code != 0:
        raise B2cError(code, where)
