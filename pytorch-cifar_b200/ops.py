"""Tensor-level wrappers over the C-ABI (include/b2c.h). torch is used only for device memory and
streams: every wrapper extracts raw pointers, passes the current CUDA stream and raises on a non-zero
return code. Activations are NHWC bf16 tensors; there is no fallback implementation here.
"""
import ctypes
import os

import torch

from . import _lib
from ._lib import BnStats, ConvDesc, RepackEntry, check

CIFAR_MEAN = (0.4914, 0.4822, 0.4465)   # reference main.py:34
CIFAR_STD = (0.2023, 0.1994, 0.2010)    # reference main.py:34
MAX_STAT_ROWS = 148          # conv epilogue partial rows (one per CTA)
MAX_DW_STAT_ROWS = 592       # depthwise fprop partial rows (one per block)


def _p(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _need(t, dtype, name):
    if t is None:
        return
    if not t.is_cuda or t.dtype != dtype or not t.is_contiguous():
        raise ValueError(f"{name}: expected contiguous CUDA {dtype}, got {t.dtype} cuda={t.is_cuda} "
                         f"contig={t.is_contiguous()}")


def conv_desc(n, h, w, cin, cout, ksize, stride, cin_valid=0, kflat_cin=0, block_n=0, max_ctas=0, flags=0,
              cout_valid=0, x_ld=0, y_ld=0):
    """x_ld / y_ld: channel pitch of the input- / output-side tensors when they are channel slices of wider ones."""
    return ConvDesc(n, h, w, cin, cout, ksize, stride, cin_valid, cout_valid, kflat_cin, block_n, max_ctas, flags,
                    x_ld, y_ld)


def pack_weight_fprop(w_oihw):
    """OIHW fp32 -> bf16 [tap][cout][cin] (torch-side helper for tests; the engine uses repack_weights)."""
    co, ci, kh, kw = w_oihw.shape
    return w_oihw.permute(2, 3, 0, 1).reshape(kh * kw, co, ci).contiguous().to(torch.bfloat16)


def pack_weight_dgrad(w_oihw):
    co, ci, kh, kw = w_oihw.shape
    return w_oihw.permute(2, 3, 1, 0).reshape(kh * kw, ci, co).contiguous().to(torch.bfloat16)


def conv_fprop(d, x, w_fprop, y, part_sum=None, part_sq=None, add_src=None):
    """y = conv(x) (+ add_src); returns the number of BatchNorm partial rows written (0 without statistics)."""
    lib = _lib.load()
    _need(x, torch.bfloat16, "x"); _need(w_fprop, torch.bfloat16, "w"); _need(y, torch.bfloat16, "y")
    _need(part_sum, torch.float32, "part_sum"); _need(part_sq, torch.float32, "part_sq")
    rows = ctypes.c_int(0)
    _need(add_src, torch.bfloat16, "add_src")
    check(lib.b2c_conv_fprop(ctypes.byref(d), _p(x), _p(w_fprop), _p(y), _p(add_src), _p(part_sum), _p(part_sq),
                             ctypes.byref(rows), _stream()), "b2c_conv_fprop")
    return rows.value


def conv1x1_cat_fprop(d, x0, x1, w_fprop2, y, part_sum=None, part_sq=None):
    """y = conv1x1(cat([x0, x1], C)): d describes one source; w_fprop2 is [2][cout][cin]. Returns the stat rows."""
    lib = _lib.load()
    _need(x0, torch.bfloat16, "x0"); _need(x1, torch.bfloat16, "x1"); _need(w_fprop2, torch.bfloat16, "w")
    _need(y, torch.bfloat16, "y"); _need(part_sum, torch.float32, "part_sum"); _need(part_sq, torch.float32, "part_sq")
    rows = ctypes.c_int(0)
    check(lib.b2c_conv1x1_cat_fprop(ctypes.byref(d), _p(x0), _p(x1), _p(w_fprop2), _p(y), _p(part_sum), _p(part_sq),
                                    ctypes.byref(rows), _stream()), "b2c_conv1x1_cat_fprop")
    return rows.value


def conv1x1_catn_fprop(d, xs, w_fprop_n, y, part_sum=None, part_sq=None):
    """y = conv1x1(cat(xs, C)) for 2..4 sources: d describes one source; w_fprop_n is [len(xs)][cout][cin]."""
    lib = _lib.load()
    for t in xs:
        _need(t, torch.bfloat16, "x")
    _need(w_fprop_n, torch.bfloat16, "w"); _need(y, torch.bfloat16, "y")
    _need(part_sum, torch.float32, "part_sum"); _need(part_sq, torch.float32, "part_sq")
    ptrs = (ctypes.c_void_p * len(xs))(*[t.data_ptr() for t in xs])
    rows = ctypes.c_int(0)
    check(lib.b2c_conv1x1_catn_fprop(ctypes.byref(d), ptrs, len(xs), _p(w_fprop_n), _p(y), _p(part_sum), _p(part_sq),
                                     ctypes.byref(rows), _stream()), "b2c_conv1x1_catn_fprop")
    return rows.value


def conv1x1_catr_fprop(d, xs, w_fprop_n, y, part_sum=None, part_sq=None):
    """y = conv1x1(cat(xs, C)) for 2..4 sources of different (padded) widths xs[t].shape[-1] <= d.cin;
    w_fprop_n is [len(xs)][cout][d.cin], slab t filled in its first xs[t].shape[-1] columns."""
    lib = _lib.load()
    for t in xs:
        _need(t, torch.bfloat16, "x")
    _need(w_fprop_n, torch.bfloat16, "w"); _need(y, torch.bfloat16, "y")
    _need(part_sum, torch.float32, "part_sum"); _need(part_sq, torch.float32, "part_sq")
    ptrs = (ctypes.c_void_p * len(xs))(*[t.data_ptr() for t in xs])
    cins = (ctypes.c_int * len(xs))(*[t.shape[-1] for t in xs])
    rows = ctypes.c_int(0)
    check(lib.b2c_conv1x1_catr_fprop(ctypes.byref(d), ptrs, cins, len(xs), _p(w_fprop_n), _p(y), _p(part_sum),
                                     _p(part_sq), ctypes.byref(rows), _stream()), "b2c_conv1x1_catr_fprop")
    return rows.value


def maxpool3x3_fwd(x, y, argmax, stride):
    lib = _lib.load()
    n, h, w, c = x.shape
    _need(x, torch.bfloat16, "x"); _need(y, torch.bfloat16, "y"); _need(argmax, torch.uint8, "argmax")
    check(lib.b2c_maxpool3x3_fwd(_p(x), _p(y), _p(argmax), n, h, w, c, stride, _stream()), "b2c_maxpool3x3_fwd")


def maxpool3x3_bwd(dy, argmax, dx, stride, add=None):
    lib = _lib.load()
    n, h, w, c = dx.shape
    _need(dy, torch.bfloat16, "dy"); _need(dx, torch.bfloat16, "dx"); _need(argmax, torch.uint8, "argmax")
    _need(add, torch.bfloat16, "add")
    check(lib.b2c_maxpool3x3_bwd(_p(dy), _p(argmax), _p(add), _p(dx), n, h, w, c, stride, _stream()),
          "b2c_maxpool3x3_bwd")


def copy_channels(src, src_off, dst, dst_off, channels):
    """dst[..., dst_off:dst_off+channels] = src[..., src_off:src_off+channels] over all rows (b2c_copy_channels)."""
    lib = _lib.load()
    _need(src, torch.bfloat16, "src"); _need(dst, torch.bfloat16, "dst")
    rows = src.numel() // src.shape[-1]
    if rows != dst.numel() // dst.shape[-1]:
        raise ValueError("copy_channels: row counts differ")
    check(lib.b2c_copy_channels(_p(src), src.shape[-1], src_off, _p(dst), dst.shape[-1], dst_off, rows, channels,
                                _stream()), "b2c_copy_channels")


def conv_dgrad(d, dy, w_dgrad, dx, add_src=None):
    lib = _lib.load()
    _need(dy, torch.bfloat16, "dy"); _need(w_dgrad, torch.bfloat16, "w"); _need(dx, torch.bfloat16, "dx")
    _need(add_src, torch.bfloat16, "add_src")
    check(lib.b2c_conv_dgrad(ctypes.byref(d), _p(dy), _p(w_dgrad), _p(dx), _p(add_src), _stream()),
          "b2c_conv_dgrad")


def conv_dgrad_bn(d, dy, w_dgrad, dz, y_bn, scale, shift, partial, add_src=None):
    """dgrad + the reduce pass of the following relu(bn(y_bn)) backward; returns the partial-row count."""
    lib = _lib.load()
    _need(dy, torch.bfloat16, "dy"); _need(w_dgrad, torch.bfloat16, "w"); _need(dz, torch.bfloat16, "dz")
    _need(y_bn, torch.bfloat16, "y_bn"); _need(add_src, torch.bfloat16, "add_src")
    _need(scale, torch.float32, "scale"); _need(shift, torch.float32, "shift"); _need(partial, torch.float32, "partial")
    rows = ctypes.c_int(0)
    check(lib.b2c_conv_dgrad_bn(ctypes.byref(d), _p(dy), _p(w_dgrad), _p(dz), _p(add_src), _p(y_bn), _p(scale),
                                _p(shift), _p(partial), ctypes.byref(rows), _stream()), "b2c_conv_dgrad_bn")
    return rows.value


def conv_dgrad_with_shortcut(d, dy, dy_shortcut, w_dgrad10, dx, add_src=None):
    """3x3 stride-2 dgrad + the 1x1 stride-2 shortcut's dgrad in one launch; w_dgrad10 = [10][cin][cout]."""
    lib = _lib.load()
    _need(dy, torch.bfloat16, "dy"); _need(dy_shortcut, torch.bfloat16, "dy_shortcut")
    _need(w_dgrad10, torch.bfloat16, "w"); _need(dx, torch.bfloat16, "dx"); _need(add_src, torch.bfloat16, "add_src")
    check(lib.b2c_conv_dgrad_with_shortcut(ctypes.byref(d), _p(dy), _p(dy_shortcut), _p(w_dgrad10), _p(dx),
                                           _p(add_src), _stream()), "b2c_conv_dgrad_with_shortcut")


WGRAD_SCRATCH_FLOATS = 148 * 5 * 128 * 64      # 24 MB: one partial tile set per CTA of the halo wgrad kernel


def conv_wgrad(d, x, dy, dw, scratch=None):
    lib = _lib.load()
    _need(x, torch.bfloat16, "x"); _need(dy, torch.bfloat16, "dy"); _need(dw, torch.float32, "dw")
    _need(scratch, torch.float32, "scratch")
    check(lib.b2c_conv_wgrad(ctypes.byref(d), _p(x), _p(dy), _p(dw), _p(scratch),
                             0 if scratch is None else scratch.numel(), _stream()), "b2c_conv_wgrad")


def dwconv_fprop(x, weight, y, c_valid, stride, part_sum=None, part_sq=None):
    """Depthwise 3x3 (groups == channels); returns the number of BatchNorm partial rows written."""
    lib = _lib.load()
    n, h, w, c = x.shape
    _need(x, torch.bfloat16, "x"); _need(y, torch.bfloat16, "y"); _need(weight, torch.float32, "weight")
    rows = ctypes.c_int(0)
    check(lib.b2c_dwconv_fprop(_p(x), _p(weight), _p(y), n, h, w, c, c_valid, stride, _p(part_sum), _p(part_sq),
                               ctypes.byref(rows), _stream()), "b2c_dwconv_fprop")
    return rows.value


def dwconv_fprop_acc(x, weight, y, c_valid, stride, stat_sum, stat_sq):
    """Depthwise 3x3 whose BatchNorm sums are accumulated into one zero-filled row (for bn_act_apply_stats)."""
    lib = _lib.load()
    n, h, w, c = x.shape
    _need(x, torch.bfloat16, "x"); _need(y, torch.bfloat16, "y"); _need(weight, torch.float32, "weight")
    _need(stat_sum, torch.float32, "sum"); _need(stat_sq, torch.float32, "sumsq")
    check(lib.b2c_dwconv_fprop_acc(_p(x), _p(weight), _p(y), n, h, w, c, c_valid, stride, _p(stat_sum), _p(stat_sq),
                                   _stream()), "b2c_dwconv_fprop_acc")


def dwconv_input_transform_ok(h, w, channels, stride):
    return _lib.load().b2c_dwconv_input_transform_ok(h, w, channels, stride) == 1


def dwconv_fprop_bnrelu(y_prev, in_scale, in_shift, weight, y, c_valid, part_sum=None, part_sq=None, accumulate=False):
    """Depthwise 3x3 stride 1 over relu(bn(y_prev)), the BatchNorm + ReLU applied while loading; returns stat rows."""
    lib = _lib.load()
    n, h, w, c = y_prev.shape
    _need(y_prev, torch.bfloat16, "y_prev"); _need(y, torch.bfloat16, "y"); _need(weight, torch.float32, "weight")
    _need(in_scale, torch.float32, "in_scale"); _need(in_shift, torch.float32, "in_shift")
    rows = ctypes.c_int(0)
    check(lib.b2c_dwconv_fprop_bnrelu(_p(y_prev), _p(in_scale), _p(in_shift), _p(weight), _p(y), n, h, w, c, c_valid,
                                      _p(part_sum), _p(part_sq), ctypes.byref(rows), 1 if accumulate else 0, _stream()),
          "b2c_dwconv_fprop_bnrelu")
    return rows.value


def dwconv_wgrad_bnrelu(y_prev, in_scale, in_shift, dy, dweight, c_valid):
    lib = _lib.load()
    n, h, w, c = y_prev.shape
    _need(dweight, torch.float32, "dweight")
    check(lib.b2c_dwconv_wgrad_bnrelu(_p(y_prev), _p(in_scale), _p(in_shift), _p(dy), _p(dweight), n, h, w, c, c_valid,
                                      _stream()), "b2c_dwconv_wgrad_bnrelu")


def dwconv_dgrad(dy, weight, dx, c_valid, stride):
    lib = _lib.load()
    n, h, w, c = dx.shape
    check(lib.b2c_dwconv_dgrad(_p(dy), _p(weight), _p(dx), n, h, w, c, c_valid, stride, _stream()), "b2c_dwconv_dgrad")


def dwconv_wgrad(x, dy, dweight, c_valid, stride):
    lib = _lib.load()
    n, h, w, c = x.shape
    _need(dweight, torch.float32, "dweight")
    check(lib.b2c_dwconv_wgrad(_p(x), _p(dy), _p(dweight), n, h, w, c, c_valid, stride, _stream()), "b2c_dwconv_wgrad")


# measurement aid: >1 issues each finalize kernel that many times (their in-graph cost = the step-time difference)
_FINALIZE_REPEAT = int(os.environ.get("B2C_EXPERIMENT_FINALIZE_REPEAT", "1"))


def bn_finalize(part_sum, part_sq, rows, channels, count, gamma, beta, eps, momentum, running_mean, running_var,
                scale, shift, save_mean, save_rstd, conv_bias=None):
    lib = _lib.load()
    for _ in range(_FINALIZE_REPEAT):
        check(lib.b2c_bn_finalize(_p(part_sum), _p(part_sq), rows, channels, float(count), _p(gamma), _p(beta),
                                  _p(conv_bias), float(eps), float(momentum), _p(running_mean), _p(running_var),
                                  _p(scale), _p(shift), _p(save_mean), _p(save_rstd), _stream()), "b2c_bn_finalize")


def bn_eval_scale_shift(channels, gamma, beta, running_mean, running_var, eps, scale, shift, conv_bias=None):
    lib = _lib.load()
    check(lib.b2c_bn_eval_scale_shift(channels, _p(gamma), _p(beta), _p(conv_bias), _p(running_mean), _p(running_var),
                                      float(eps),
                                      _p(scale), _p(shift), _stream()), "b2c_bn_eval_scale_shift")


def bn_act_apply(y, scale, shift, out, rows, channels, relu=True, res=None, res_scale=None, res_shift=None,
                 relu_bits=None):
    """relu_bits: optional uint8 [rows*channels/8] receiving the ReLU mask (one bit per element)."""
    lib = _lib.load()
    _need(relu_bits, torch.uint8, "relu_bits")
    check(lib.b2c_bn_act_apply(_p(y), _p(scale), _p(shift), _p(res), _p(res_scale), _p(res_shift), _p(out),
                               rows, channels, 1 if relu else 0, _p(relu_bits), _stream()), "b2c_bn_act_apply")


def bn_stats(stat_sum, stat_sq, gamma, beta, count, eps, momentum, running_mean, running_var, scale, shift, save_mean,
             save_rstd, conv_bias=None):
    """B2cBnStats record for bn_act_apply_stats (keep the tensors alive for as long as the record is used)."""
    def a(t):
        return None if t is None else t.data_ptr()
    return BnStats(a(stat_sum), a(stat_sq), a(gamma), a(beta), a(conv_bias), a(running_mean), a(running_var), a(scale),
                   a(shift), a(save_mean), a(save_rstd), float(count), float(eps), float(momentum))


def bn_act_apply_stats(y, bn, out, rows, channels, relu=True, res=None, res_bn=None, relu_bits=None):
    """bn_finalize + bn_act_apply in one launch from accumulated statistics (bn / res_bn: ops.bn_stats records)."""
    lib = _lib.load()
    _need(relu_bits, torch.uint8, "relu_bits")
    check(lib.b2c_bn_act_apply_stats(_p(y), ctypes.byref(bn), _p(res), None if res_bn is None else ctypes.byref(res_bn),
                                     _p(out), rows, channels, 1 if relu else 0, _p(relu_bits), _stream()),
          "b2c_bn_act_apply_stats")


def bn_bwd_reduce_acc(dout, act, y, dz_out, rows, channels, sums, relu_scale=None, relu_shift=None, relu_bits=None):
    lib = _lib.load()
    _need(relu_bits, torch.uint8, "relu_bits"); _need(sums, torch.float32, "sums")
    check(lib.b2c_bn_bwd_reduce_acc(_p(dout), _p(act), _p(y), _p(dz_out), rows, channels, _p(sums), _p(relu_scale),
                                    _p(relu_shift), _p(relu_bits), _stream()), "b2c_bn_bwd_reduce_acc")


def bn_bwd_apply_stats(dz, y, sums, count, gamma, save_mean, save_rstd, dgamma, dbeta, dy, rows, channels, add=None,
                       relu_scale=None, relu_shift=None):
    """bn_bwd_finalize + bn_bwd_apply in one launch from one accumulated row sums[2][channels]."""
    lib = _lib.load()
    check(lib.b2c_bn_bwd_apply_stats(_p(dz), _p(y), _p(sums), float(count), _p(gamma), _p(save_mean), _p(save_rstd),
                                     _p(dgamma), _p(dbeta), _p(add), _p(dy), rows, channels, _p(relu_scale),
                                     _p(relu_shift), _stream()), "b2c_bn_bwd_apply_stats")


def bn_bwd_partial_rows(rows, channels):
    lib = _lib.load()
    r = lib.b2c_bn_bwd_partial_rows(rows, channels)
    if r < 0:
        check(r, "b2c_bn_bwd_partial_rows")
    return r


def bn_bwd_reduce(dout, act, y, dz_out, rows, channels, partial, relu_scale=None, relu_shift=None, relu_bits=None):
    lib = _lib.load()
    _need(relu_bits, torch.uint8, "relu_bits")
    check(lib.b2c_bn_bwd_reduce(_p(dout), _p(act), _p(y), _p(dz_out), rows, channels, _p(partial), _p(relu_scale),
                                _p(relu_shift), _p(relu_bits), _stream()), "b2c_bn_bwd_reduce")


def bn_bwd_finalize(partial, partial_rows, channels, count, gamma, save_mean, save_rstd, dgamma, dbeta, ca, cb, cc):
    lib = _lib.load()
    for _ in range(_FINALIZE_REPEAT):
        check(lib.b2c_bn_bwd_finalize(_p(partial), partial_rows, channels, float(count), _p(gamma), _p(save_mean),
                                      _p(save_rstd), _p(dgamma), _p(dbeta), _p(ca), _p(cb), _p(cc), _stream()),
              "b2c_bn_bwd_finalize")


NOALLOC_MAX_CHANNELS = 512


def bn_bwd_stats_noalloc(dout, act, y, dz_out, rows, channels, count, gamma, save_mean, save_rstd, dgamma, dbeta,
                         ca, cb, cc):
    """bn_bwd_reduce + bn_bwd_finalize without any scratch argument: for callers that cannot allocate device
    memory around the call (see include/b2c.h). channels <= NOALLOC_MAX_CHANNELS; calls must not overlap."""
    lib = _lib.load()
    check(lib.b2c_bn_bwd_stats_noalloc(_p(dout), _p(act), _p(y), _p(dz_out), rows, channels, float(count), _p(gamma),
                                       _p(save_mean), _p(save_rstd), _p(dgamma), _p(dbeta), _p(ca), _p(cb), _p(cc),
                                       _stream()), "b2c_bn_bwd_stats_noalloc")


def bn_bwd_apply(dz, y, ca, cb, cc, dy, rows, channels, add=None, relu_scale=None, relu_shift=None):
    lib = _lib.load()
    check(lib.b2c_bn_bwd_apply(_p(dz), _p(y), _p(ca), _p(cb), _p(cc), _p(add), _p(dy), rows, channels,
                               _p(relu_scale), _p(relu_shift), _stream()), "b2c_bn_bwd_apply")


def maxpool2x2_fwd(x, y):
    lib = _lib.load()
    n, h, w, c = x.shape
    _need(x, torch.bfloat16, "x"); _need(y, torch.bfloat16, "y")
    check(lib.b2c_maxpool2x2_fwd(_p(x), _p(y), n, h, w, c, _stream()), "b2c_maxpool2x2_fwd")


def maxpool2x2_bwd(x, y, dy, dx):
    lib = _lib.load()
    n, h, w, c = x.shape
    check(lib.b2c_maxpool2x2_bwd(_p(x), _p(y), _p(dy), _p(dx), n, h, w, c, _stream()), "b2c_maxpool2x2_bwd")


def head_fwd(act, n, hw, channels, weight, bias, classes, pooled, logits, act_ld=None):
    """act_ld: channel pitch of act when it exceeds `channels` (padded layouts) -> b2c_head_fwd_ld."""
    lib = _lib.load()
    if act_ld is not None and act_ld != channels:
        check(lib.b2c_head_fwd_ld(_p(act), n, hw, channels, act_ld, _p(weight), _p(bias), classes, _p(pooled),
                                  _p(logits), _stream()), "b2c_head_fwd_ld")
        return
    check(lib.b2c_head_fwd(_p(act), n, hw, channels, _p(weight), _p(bias), classes, _p(pooled), _p(logits),
                           _stream()), "b2c_head_fwd")


def ce_loss(logits, target, n, classes, dlogits, loss, correct=None):
    lib = _lib.load()
    _need(target, torch.int64, "target")
    check(lib.b2c_ce_loss(_p(logits), _p(target), n, classes, _p(dlogits), _p(loss), _p(correct), _stream()),
          "b2c_ce_loss")


def head_bwd(dlogits, pooled, weight, n, hw, channels, classes, dact, dweight, dbias, act_ld=None):
    lib = _lib.load()
    if act_ld is not None and act_ld != channels:
        check(lib.b2c_head_bwd_ld(_p(dlogits), _p(pooled), _p(weight), n, hw, channels, act_ld, classes, _p(dact),
                                  _p(dweight), _p(dbias), _stream()), "b2c_head_bwd_ld")
        return
    check(lib.b2c_head_bwd(_p(dlogits), _p(pooled), _p(weight), n, hw, channels, classes, _p(dact), _p(dweight),
                           _p(dbias), _stream()), "b2c_head_bwd")


def se_fwd(y, n, hw, channels, ld, reduced, w1, b1, w2, b2, pooled, hidden, scale, out, add=None):
    """out = y * sigmoid(fc2(relu(fc1(avg_pool(y))))) (+ add) — b2c_se_fwd; w1 [reduced, channels(,1,1)] etc."""
    lib = _lib.load()
    _need(y, torch.bfloat16, "y"); _need(out, torch.bfloat16, "out"); _need(add, torch.bfloat16, "add")
    for t, name in ((w1, "w1"), (b1, "b1"), (w2, "w2"), (b2, "b2"), (pooled, "pooled"), (hidden, "hidden"), (scale, "scale")):
        _need(t, torch.float32, name)
    check(lib.b2c_se_fwd(_p(y), n, hw, channels, ld, reduced, _p(w1), _p(b1), _p(w2), _p(b2), _p(pooled), _p(hidden),
                         _p(scale), _p(add), _p(out), _stream()), "b2c_se_fwd")


def se_bwd(dout, y, n, hw, channels, ld, reduced, w1, w2, pooled, hidden, scale, ds, dz1, dp, dw1, db1, dw2, db2, dy):
    lib = _lib.load()
    _need(dout, torch.bfloat16, "dout"); _need(y, torch.bfloat16, "y"); _need(dy, torch.bfloat16, "dy")
    for t, name in ((w1, "w1"), (w2, "w2"), (pooled, "pooled"), (hidden, "hidden"), (scale, "scale"), (ds, "ds"),
                    (dz1, "dz1"), (dp, "dp"), (dw1, "dw1"), (db1, "db1"), (dw2, "dw2"), (db2, "db2")):
        _need(t, torch.float32, name)
    check(lib.b2c_se_bwd(_p(dout), _p(y), n, hw, channels, ld, reduced, _p(w1), _p(w2), _p(pooled), _p(hidden),
                         _p(scale), _p(ds), _p(dz1), _p(dp), _p(dw1), _p(db1), _p(dw2), _p(db2), _p(dy), _stream()),
          "b2c_se_bwd")


_F3 = ctypes.c_float * 3


def augment_nchw(data_u8, idx, ijf, out, mean=CIFAR_MEAN, std=CIFAR_STD):
    lib = _lib.load()
    _need(data_u8, torch.uint8, "data"); _need(idx, torch.int64, "idx"); _need(ijf, torch.int8, "ijf")
    _need(out, torch.float32, "out")
    check(lib.b2c_augment_nchw(_p(data_u8), _p(idx), _p(ijf), _p(out), idx.numel(), _F3(*mean), _F3(*std),
                               _stream()), "b2c_augment_nchw")


def augment_im2col(data_u8, idx, ijf, out, mean=CIFAR_MEAN, std=CIFAR_STD, taps=9):
    """taps = 9: im2col operand of a 3x3 stem; taps = 1: the image padded to 64 channels (1x1 stem)."""
    lib = _lib.load()
    _need(data_u8, torch.uint8, "data"); _need(idx, torch.int64, "idx"); _need(ijf, torch.int8, "ijf")
    _need(out, torch.bfloat16, "out")
    fn, name = (lib.b2c_augment_im2col, "b2c_augment_im2col") if taps == 9 else \
        (lib.b2c_augment_nhwc64, "b2c_augment_nhwc64")
    check(fn(_p(data_u8), _p(idx), _p(ijf), _p(out), idx.numel(), _F3(*mean), _F3(*std), _stream()), name)


def stem_im2col(x_nchw, out, taps=9):
    lib = _lib.load()
    _need(x_nchw, torch.float32, "x"); _need(out, torch.bfloat16, "out")
    if taps == 9:
        check(lib.b2c_stem_im2col(_p(x_nchw), _p(out), x_nchw.shape[0], _stream()), "b2c_stem_im2col")
    else:
        check(lib.b2c_stem_nhwc64(_p(x_nchw), _p(out), x_nchw.shape[0], _stream()), "b2c_stem_nhwc64")


def grouped_wgrad_extract(dense_tco, dw_oihw, cout, cin, taps, groups):
    """dw_oihw[o][ci][tap] += dense_tco[tap][(o // (cout/g)) * (cin/g) + ci][o] (b2c_grouped_wgrad_extract)."""
    lib = _lib.load()
    _need(dense_tco, torch.float32, "dense"); _need(dw_oihw, torch.float32, "dw")
    if dense_tco.numel() < taps * cin * cout or dw_oihw.numel() != cout * (cin // groups) * taps:
        raise ValueError("grouped_wgrad_extract: buffer sizes do not match the conv")
    check(lib.b2c_grouped_wgrad_extract(_p(dense_tco), _p(dw_oihw), cout, cin, taps, groups, _stream()),
          "b2c_grouped_wgrad_extract")


def sgd_step(param, grad, buf, hyper):
    lib = _lib.load()
    _need(param, torch.float32, "param"); _need(grad, torch.float32, "grad"); _need(buf, torch.float32, "buf")
    _need(hyper, torch.float32, "hyper")
    check(lib.b2c_sgd_step(_p(param), _p(grad), _p(buf), param.numel(), _p(hyper), _stream()), "b2c_sgd_step")


def make_repack_table(entries, device):
    """entries: list of dicts with the B2cRepackEntry fields -> device uint8 tensor holding the table."""
    arr = (RepackEntry * len(entries))()
    for i, e in enumerate(entries):
        arr[i] = RepackEntry(e["src_off"], e["wf_off"], e.get("wd_off", -1), e["cout"], e["cin"], e["taps"],
                             e["cin_pad"], e.get("kflat", 0), e.get("cout_pad", 0))
    raw = bytes(arr)
    t = torch.frombuffer(bytearray(raw), dtype=torch.uint8).clone()
    return t.to(device)


def repack_weights(master, wf_arena, wd_arena, table_dev, entries):
    lib = _lib.load()
    check(lib.b2c_repack_weights(_p(master), _p(wf_arena), _p(wd_arena), _p(table_dev), entries, _stream()),
          "b2c_repack_weights")
