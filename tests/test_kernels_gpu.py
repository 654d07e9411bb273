"""GPU parity tests of each C-ABI kernel against a plain PyTorch fp32 restatement of the same op.

Every call goes through the C-ABI (ctypes -> libb2c.so). Tolerances: bf16-stored results within 1e-2
relative (Frobenius) of the fp32 reference (BASELINE.json north_star); fp32 results within 1e-3.
"""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

BF16_TOL = 1e-2
F32_TOL = 1e-3


@pytest.fixture(scope="module")
def ops():
    import b200cifar
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    return b200cifar.ops


def rel_err(a, b):
    a = a.double()
    b = b.double()
    return ((a - b).norm() / b.norm().clamp_min(1e-30)).item()


def nhwc_bf16(n, h, w, c, seed, scale=1.0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(n, h, w, c, device="cuda", generator=g) * scale).to(torch.bfloat16).contiguous()


def weight(cout, cin, k, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    w = torch.randn(cout, cin, k, k, device="cuda", generator=g) * (2.0 / (cin * k * k)) ** 0.5
    return w.to(torch.bfloat16).float()  # values exactly representable in bf16


def to_nchw(x):
    return x.float().permute(0, 3, 1, 2).contiguous()


CONV_SHAPES = [
    # n, h, w, cin, cout, k, stride
    (4, 32, 32, 64, 64, 3, 1),
    (8, 16, 16, 128, 128, 3, 1),
    (16, 8, 8, 256, 256, 3, 1),
    (32, 4, 4, 512, 512, 3, 1),
    (4, 32, 32, 64, 128, 3, 2),
    (8, 16, 16, 128, 256, 3, 2),
    (16, 8, 8, 256, 512, 3, 2),
    (4, 32, 32, 64, 128, 1, 2),
    (16, 8, 8, 256, 512, 1, 2),
    (2, 32, 32, 64, 64, 1, 1),
    (5, 4, 4, 512, 512, 3, 1),     # ragged: 80 pixels, one partial tile
    (3, 8, 8, 128, 128, 3, 1),     # ragged: 192 pixels
    (64, 2, 2, 512, 512, 3, 1),    # VGG last stage
]


@pytest.mark.parametrize("shape", CONV_SHAPES)
def test_conv_fprop(ops, shape):
    n, h, w, cin, cout, k, stride = shape
    x = nhwc_bf16(n, h, w, cin, 1)
    wt = weight(cout, cin, k, 2)
    ho, wo = h // stride, w // stride
    y = torch.full((n, ho, wo, cout), float("nan"), device="cuda", dtype=torch.bfloat16)
    ps = torch.zeros(ops.MAX_STAT_ROWS, cout, device="cuda")
    pq = torch.zeros(ops.MAX_STAT_ROWS, cout, device="cuda")
    d = ops.conv_desc(n, h, w, cin, cout, k, stride)
    rows = ops.conv_fprop(d, x, ops.pack_weight_fprop(wt), y, ps, pq)
    torch.cuda.synchronize()
    ref = F.conv2d(to_nchw(x), wt, stride=stride, padding=k // 2)
    got = to_nchw(y)
    assert torch.isfinite(got).all()
    assert rel_err(got, ref) < BF16_TOL
    assert 1 <= rows <= ops.MAX_STAT_ROWS
    s1 = ps[:rows].sum(0)
    s2 = pq[:rows].sum(0)
    assert rel_err(s1, ref.sum((0, 2, 3))) < 2e-3 or (s1 - ref.sum((0, 2, 3))).abs().max() < 1e-1
    assert rel_err(s2, (ref * ref).sum((0, 2, 3))) < F32_TOL


# The batch bench.py is quoted on (BASELINE config 2 on one GPU: 1024 images): tile counts, persistent-loop wave
# counts, wgrad split-K depth (K = 1 048 576 pixels on layer 1) and halo-wgrad scratch use all differ from the
# small cases above, so the same three kernels are checked at that size with the flags the engine passes.
BENCH_SHAPES = [
    (1024, 32, 32, 64, 64, 3, 1),      # layer1 conv (conv_halo2 RES, wgrad_halo)
    (1024, 32, 32, 64, 128, 3, 2),     # layer2.0.conv1
    (1024, 32, 32, 64, 128, 1, 2),     # layer2.0.shortcut
    (1024, 16, 16, 128, 128, 3, 1),    # layer2 conv (conv_halo2 BN=128, wgrad_row)
    (1024, 8, 8, 256, 256, 3, 1),      # layer3 conv (MODE 1 tiles, wgrad_row2)
    (1024, 4, 4, 512, 512, 3, 1),      # layer4 conv (tap_gemm2 BN=256)
]


@pytest.mark.parametrize("shape", BENCH_SHAPES)
def test_conv_at_bench_batch(ops, shape):
    n, h, w, cin, cout, k, stride = shape
    ho, wo = h // stride, w // stride
    x = nhwc_bf16(n, h, w, cin, 71)
    dy = nhwc_bf16(n, ho, wo, cout, 72, scale=0.1)
    wt = weight(cout, cin, k, 73)
    xf, dyf = to_nchw(x), to_nchw(dy)
    pair = 128
    d = ops.conv_desc(n, h, w, cin, cout, k, stride, flags=pair)
    # fprop + BatchNorm partial sums
    y = torch.full((n, ho, wo, cout), float("nan"), device="cuda", dtype=torch.bfloat16)
    ps = torch.zeros(ops.MAX_STAT_ROWS, cout, device="cuda")
    pq = torch.zeros(ops.MAX_STAT_ROWS, cout, device="cuda")
    rows = ops.conv_fprop(d, x, ops.pack_weight_fprop(wt), y, ps, pq)
    torch.cuda.synchronize()
    ref = F.conv2d(xf, wt, stride=stride, padding=k // 2)
    assert torch.isfinite(to_nchw(y)).all() and rel_err(to_nchw(y), ref) < BF16_TOL
    assert rel_err(pq[:rows].sum(0), (ref * ref).sum((0, 2, 3))) < F32_TOL
    assert (ps[:rows].sum(0) - ref.sum((0, 2, 3))).abs().max() < 2e-3 * ref.abs().sum((0, 2, 3)).max()
    del ref
    # dgrad
    dx = torch.full((n, h, w, cin), float("nan"), device="cuda", dtype=torch.bfloat16)
    if k == 1 and stride == 2:
        ops.conv_dgrad(d, dy, ops.pack_weight_dgrad(wt), dx)
    else:
        add = nhwc_bf16(n, h, w, cin, 74, scale=0.1)
        ops.conv_dgrad(d, dy, ops.pack_weight_dgrad(wt), dx, add)
    torch.cuda.synchronize()
    ref = torch.nn.grad.conv2d_input((n, cin, h, w), wt, dyf, stride=stride, padding=k // 2)
    if not (k == 1 and stride == 2):
        ref = ref + to_nchw(add)
    assert torch.isfinite(to_nchw(dx)).all() and rel_err(to_nchw(dx), ref) < BF16_TOL
    del ref
    # wgrad, [tap][cin][cout] layout + scratch, as the engine calls it
    refw = torch.nn.grad.conv2d_weight(xf, (cout, cin, k, k), dyf, stride=stride, padding=k // 2)
    scratch = torch.zeros(ops.WGRAD_SCRATCH_FLOATS, device="cuda")
    dw = torch.zeros(k * k, cin, cout, device="cuda")
    ops.conv_wgrad(ops.conv_desc(n, h, w, cin, cout, k, stride, flags=32 | pair), x, dy, dw, scratch=scratch)
    torch.cuda.synchronize()
    assert rel_err(dw.view(k, k, cin, cout).permute(3, 2, 0, 1), refw) < 2e-3


@pytest.mark.parametrize("shape", [s for s in CONV_SHAPES if s[5] == 3 and s[6] == 1 and s[1] in (8, 16, 32)])
def test_conv_per_tap_kernel_still_covered(ops, shape):
    """flags=1 forces the per-tap kernel on shapes that normally take the halo-reuse kernel."""
    n, h, w, cin, cout, k, stride = shape
    x = nhwc_bf16(n, h, w, cin, 41)
    wt = weight(cout, cin, k, 42)
    y = torch.empty(n, h, w, cout, device="cuda", dtype=torch.bfloat16)
    d = ops.conv_desc(n, h, w, cin, cout, k, stride, flags=1)
    ops.conv_fprop(d, x, ops.pack_weight_fprop(wt), y)
    dy = nhwc_bf16(n, h, w, cout, 43)
    dx = torch.empty(n, h, w, cin, device="cuda", dtype=torch.bfloat16)
    ops.conv_dgrad(d, dy, ops.pack_weight_dgrad(wt), dx)
    torch.cuda.synchronize()
    assert rel_err(to_nchw(y), F.conv2d(to_nchw(x), wt, padding=1)) < BF16_TOL
    ref = torch.nn.grad.conv2d_input((n, cin, h, w), wt, to_nchw(dy), padding=1)
    assert rel_err(to_nchw(dx), ref) < BF16_TOL


@pytest.mark.parametrize("bn", [64, 128, 256])
def test_conv_fprop_block_n(ops, bn):
    n, h, w, cin, cout, k, stride = 8, 16, 16, 128, 256, 3, 1
    x = nhwc_bf16(n, h, w, cin, 3)
    wt = weight(cout, cin, k, 4)
    y = torch.empty(n, h, w, cout, device="cuda", dtype=torch.bfloat16)
    d = ops.conv_desc(n, h, w, cin, cout, k, stride, block_n=bn)
    ops.conv_fprop(d, x, ops.pack_weight_fprop(wt), y)
    torch.cuda.synchronize()
    ref = F.conv2d(to_nchw(x), wt, stride=stride, padding=1)
    assert rel_err(to_nchw(y), ref) < BF16_TOL


def test_conv_fprop_multi_wave(ops):
    # more tiles than CTAs so the persistent loop, the smem ring wrap and both TMEM buffers are used
    n, h, w, cin, cout = 64, 32, 32, 64, 64
    x = nhwc_bf16(n, h, w, cin, 5)
    wt = weight(cout, cin, 3, 6)
    y = torch.empty(n, h, w, cout, device="cuda", dtype=torch.bfloat16)
    ps = torch.zeros(ops.MAX_STAT_ROWS, cout, device="cuda")
    pq = torch.zeros(ops.MAX_STAT_ROWS, cout, device="cuda")
    d = ops.conv_desc(n, h, w, cin, cout, 3, 1, max_ctas=37)
    rows = ops.conv_fprop(d, x, ops.pack_weight_fprop(wt), y, ps, pq)
    torch.cuda.synchronize()
    ref = F.conv2d(to_nchw(x), wt, padding=1)
    assert rel_err(to_nchw(y), ref) < BF16_TOL
    assert rel_err(pq[:rows].sum(0), (ref * ref).sum((0, 2, 3))) < F32_TOL


@pytest.mark.parametrize("shape", CONV_SHAPES)
@pytest.mark.parametrize("with_add", [False, True])
def test_conv_dgrad(ops, shape, with_add):
    n, h, w, cin, cout, k, stride = shape
    ho, wo = h // stride, w // stride
    dy = nhwc_bf16(n, ho, wo, cout, 7)
    wt = weight(cout, cin, k, 8)
    ref = torch.nn.grad.conv2d_input((n, cin, h, w), wt, to_nchw(dy), stride=stride, padding=k // 2)
    dx = torch.full((n, h, w, cin), float("nan"), device="cuda", dtype=torch.bfloat16)
    add = None
    if with_add:
        if k == 1 and stride == 2:
            # in-place accumulate form
            dx = nhwc_bf16(n, h, w, cin, 9)
            ref = ref + to_nchw(dx)
            add = dx
        else:
            add = nhwc_bf16(n, h, w, cin, 9)
            ref = ref + to_nchw(add)
    d = ops.conv_desc(n, h, w, cin, cout, k, stride)
    ops.conv_dgrad(d, dy, ops.pack_weight_dgrad(wt), dx, add)
    torch.cuda.synchronize()
    got = to_nchw(dx)
    assert torch.isfinite(got).all()
    assert rel_err(got, ref) < BF16_TOL


@pytest.mark.parametrize("n,h,cin,cout", [(8, 32, 64, 128), (16, 16, 128, 256), (64, 8, 256, 512), (3, 16, 64, 64)])
@pytest.mark.parametrize("with_add", [False, True])
@pytest.mark.parametrize("pair", [0, 128])
def test_conv_dgrad_with_shortcut(ops, n, h, cin, cout, with_add, pair):
    """3x3 stride-2 dgrad + 1x1 stride-2 shortcut dgrad as one launch (four output phases as tiles) against the sum
    of the two torch input gradients; and the same 3x3 dgrad alone must agree with the traversal-free path."""
    ho = h // 2
    dy, dys = nhwc_bf16(n, ho, ho, cout, 31), nhwc_bf16(n, ho, ho, cout, 32)
    w3, w1 = weight(cout, cin, 3, 33), weight(cout, cin, 1, 34)
    ref = (torch.nn.grad.conv2d_input((n, cin, h, h), w3, to_nchw(dy), stride=2, padding=1) +
           torch.nn.grad.conv2d_input((n, cin, h, h), w1, to_nchw(dys), stride=2, padding=0))
    add = None
    if with_add:
        add = nhwc_bf16(n, h, h, cin, 35)
        ref = ref + to_nchw(add)
    w10 = torch.cat([ops.pack_weight_dgrad(w3), ops.pack_weight_dgrad(w1)], 0).contiguous()
    assert w10.shape == (10, cin, cout)
    dx = torch.full((n, h, h, cin), float("nan"), device="cuda", dtype=torch.bfloat16)
    d = ops.conv_desc(n, h, h, cin, cout, 3, 2, flags=pair)      # flags bit 7: the CTA-pair kernel
    ops.conv_dgrad_with_shortcut(d, dy, dys, w10, dx, add)
    torch.cuda.synchronize()
    got = to_nchw(dx)
    assert torch.isfinite(got).all()
    assert rel_err(got, ref) < BF16_TOL
    with pytest.raises(RuntimeError):
        ops.conv_dgrad_with_shortcut(ops.conv_desc(n, h, h, cin, cout, 3, 1), dy, dys, w10, dx, add)


@pytest.mark.parametrize("shape", [(4, 32, 32, 64, 64, 3, 1), (8, 16, 16, 128, 128, 3, 1), (16, 8, 8, 256, 256, 3, 1),
                                   (32, 4, 4, 512, 512, 3, 1), (8, 16, 16, 64, 128, 3, 2), (4, 16, 16, 64, 256, 1, 1),
                                   (3, 8, 8, 128, 128, 3, 1), (5, 4, 4, 64, 64, 3, 1)])
def test_conv_dgrad_bn_fused(ops, shape):
    """b2c_conv_dgrad_bn == b2c_conv_dgrad followed by b2c_bn_bwd_reduce with the mask taken from y: the masked
    gradient must be bit-identical, the partial sums equal up to fp32 summation order (and the fused sums see
    the gradient before its bf16 rounding)."""
    n, h, w, cin, cout, k, stride = shape
    ho, wo = h // stride, w // stride
    dy = nhwc_bf16(n, ho, wo, cout, 51)
    wd = ops.pack_weight_dgrad(weight(cout, cin, k, 52))
    y_bn = nhwc_bf16(n, h, w, cin, 53)
    g = torch.Generator(device="cuda").manual_seed(54)
    scale = torch.rand(cin, device="cuda", generator=g) + 0.5
    shift = torch.randn(cin, device="cuda", generator=g) * 0.3
    d = ops.conv_desc(n, h, w, cin, cout, k, stride)
    rows = n * h * w
    # two-pass reference through the unfused entry points
    dx = torch.empty(n, h, w, cin, device="cuda", dtype=torch.bfloat16)
    ops.conv_dgrad(d, dy, wd, dx)
    prow_ref = ops.bn_bwd_partial_rows(rows, cin)
    part_ref = torch.zeros(prow_ref, 2, cin, device="cuda")
    ops.bn_bwd_reduce(dx, None, y_bn, None, rows, cin, part_ref, scale, shift)
    z = y_bn.double() * scale.double() + shift.double()
    eps = 1e-6 * (y_bn.double().abs() * scale.double().abs() + shift.double().abs())
    # fused
    dz = torch.full_like(dx, float("nan"))
    part = torch.full((148, 2, cin), float("nan"), device="cuda")
    prow = ops.conv_dgrad_bn(d, dy, wd, dz, y_bn, scale, shift, part)
    torch.cuda.synchronize()
    assert 1 <= prow <= 148
    assert torch.isfinite(dz.float()).all() and torch.isfinite(part[:prow]).all()
    on, off = z > eps, z < -eps                     # (the few elements within rounding of 0 may go either way)
    assert torch.equal(dz[on], dx[on]) and (dz[off] == 0).all()
    assert (on | off).float().mean() > 0.999
    got, ref = part[:prow].double().sum(0), part_ref.double().sum(0)
    scale_ = ref.abs().max()
    assert (got - ref).abs().max() <= 2e-3 * scale_ + 1e-3, ((got - ref).abs().max().item(), scale_.item())


@pytest.mark.parametrize("shape", [(8, 32, 32, 64, 128, 3, 2), (8, 16, 16, 128, 256, 1, 2)])
def test_stride2_split_w_map_equals_traversal_stride(ops, shape):
    """flags bit 2 selects the TMA traversal-stride addressing of x; the default split-W map must give the same
    bits (same operands, same MMA order)."""
    n, h, w, cin, cout, k, stride = shape
    x = nhwc_bf16(n, h, w, cin, 41)
    wt = weight(cout, cin, k, 42)
    ys = []
    for flags in (0, 4):
        y = torch.empty(n, h // 2, w // 2, cout, device="cuda", dtype=torch.bfloat16)
        ops.conv_fprop(ops.conv_desc(n, h, w, cin, cout, k, stride, flags=flags), x, ops.pack_weight_fprop(wt), y)
        ys.append(y)
    torch.cuda.synchronize()
    assert torch.equal(ys[0], ys[1])


@pytest.mark.parametrize("shape", [(4, 32, 32, 64, 64, 3, 1), (8, 16, 16, 128, 128, 3, 1), (16, 8, 8, 256, 512, 3, 2),
                                   (32, 4, 4, 512, 512, 3, 1), (8, 16, 16, 128, 256, 1, 2), (3, 8, 8, 128, 128, 3, 1),
                                   (4, 16, 16, 64, 128, 1, 1), (16, 8, 8, 256, 256, 3, 1), (2, 16, 16, 128, 256, 3, 1),
                                   (40, 16, 16, 256, 128, 3, 1), (300, 8, 8, 128, 128, 3, 1)])
def test_conv_wgrad_tap_cin_cout_layout(ops, shape):
    """flags bit 5: the weight gradient lands in [tap][cin][cout] order (vector reductions, 4 rows x 128 B per
    instruction) and must equal the OIHW result permuted; it accumulates on top of existing contents."""
    n, h, w, cin, cout, k, stride = shape
    ho, wo = h // stride, w // stride
    x = nhwc_bf16(n, h, w, cin, 61)
    dy = nhwc_bf16(n, ho, wo, cout, 62)
    ref = torch.nn.grad.conv2d_weight(to_nchw(x), (cout, cin, k, k), to_nchw(dy), stride=stride, padding=k // 2)
    scratch = torch.zeros(ops.WGRAD_SCRATCH_FLOATS, device="cuda")
    for use_scratch in (False, True):
        dw = torch.full((k * k, cin, cout), 0.5, device="cuda")
        d = ops.conv_desc(n, h, w, cin, cout, k, stride, flags=32)
        ops.conv_wgrad(d, x, dy, dw, scratch=scratch if use_scratch else None)
        torch.cuda.synchronize()
        got = (dw - 0.5).view(k, k, cin, cout).permute(3, 2, 0, 1)
        assert rel_err(got, ref) < 2e-3
    # the row kernel (16x16 / 8x8 images, 128-channel units) against the per-tap kernel (flags bit 6)
    dw2 = torch.zeros(k * k, cin, cout, device="cuda")
    ops.conv_wgrad(ops.conv_desc(n, h, w, cin, cout, k, stride, flags=32 | 64), x, dy, dw2)
    torch.cuda.synchronize()
    assert rel_err(dw2.view(k, k, cin, cout).permute(3, 2, 0, 1), ref) < 2e-3
    # ... and the CTA-pair row kernel (flags bit 7; odd unit counts leave the last pair's peer without rows)
    dw3 = torch.zeros(k * k, cin, cout, device="cuda")
    ops.conv_wgrad(ops.conv_desc(n, h, w, cin, cout, k, stride, flags=32 | 128 | 256), x, dy, dw3)
    torch.cuda.synchronize()
    assert rel_err(dw3.view(k, k, cin, cout).permute(3, 2, 0, 1), ref) < 2e-3


def test_repack_from_tap_cin_cout_master(ops):
    """B2cRepackEntry.kflat == 2: master weights stored [tap][cin][cout] give the same two bf16 operands as the
    OIHW master."""
    g = torch.Generator(device="cuda").manual_seed(63)
    entries, masters, offs = [], [], 0
    wf_n = wd_n = 0
    shapes = [(64, 64, 3), (128, 64, 1), (96, 24, 1), (512, 256, 3)]
    for cout, cin, k in shapes:
        w = torch.randn(cout, cin, k, k, device="cuda", generator=g)
        masters.append(w)
        cin_p, cout_p = (cin + 63) // 64 * 64, (cout + 63) // 64 * 64
        entries.append(dict(src_off=offs, wf_off=wf_n, wd_off=wd_n, cout=cout, cin=cin, taps=k * k, cin_pad=cin_p,
                            cout_pad=cout_p, kflat=2))
        offs += w.numel(); wf_n += cout_p * cin_p * k * k; wd_n += cout_p * cin_p * k * k
    master = torch.cat([w.permute(2, 3, 1, 0).reshape(-1) for w in masters]).contiguous()
    wf = torch.zeros(wf_n, device="cuda", dtype=torch.bfloat16)
    wd = torch.zeros(wd_n, device="cuda", dtype=torch.bfloat16)
    ops.repack_weights(master, wf, wd, ops.make_repack_table(entries, "cuda"), len(entries))
    torch.cuda.synchronize()
    for e, w in zip(entries, masters):
        cout, cin, kk = e["cout"], e["cin"], e["taps"]
        f = wf[e["wf_off"]:e["wf_off"] + kk * e["cout_pad"] * e["cin_pad"]].view(kk, e["cout_pad"], e["cin_pad"])
        d = wd[e["wd_off"]:e["wd_off"] + kk * e["cout_pad"] * e["cin_pad"]].view(kk, e["cin_pad"], e["cout_pad"])
        assert torch.equal(f[:, :cout, :cin], ops.pack_weight_fprop(w))
        assert torch.equal(d[:, :cin, :cout], ops.pack_weight_dgrad(w))
        assert f[:, cout:].abs().sum() == 0 and f[:, :, cin:].abs().sum() == 0
        assert d[:, cin:].abs().sum() == 0 and d[:, :, cout:].abs().sum() == 0


@pytest.mark.parametrize("shape", CONV_SHAPES)
def test_conv_wgrad(ops, shape):
    n, h, w, cin, cout, k, stride = shape
    ho, wo = h // stride, w // stride
    x = nhwc_bf16(n, h, w, cin, 10)
    dy = nhwc_bf16(n, ho, wo, cout, 11)
    ref = torch.nn.grad.conv2d_weight(to_nchw(x), (cout, cin, k, k), to_nchw(dy), stride=stride, padding=k // 2)
    dw = torch.zeros(cout, cin, k, k, device="cuda")
    d = ops.conv_desc(n, h, w, cin, cout, k, stride)
    ops.conv_wgrad(d, x, dy, dw)
    torch.cuda.synchronize()
    assert rel_err(dw, ref) < F32_TOL
    # accumulation semantics: a second call doubles the result
    ops.conv_wgrad(d, x, dy, dw)
    torch.cuda.synchronize()
    assert rel_err(dw, 2 * ref) < F32_TOL


@pytest.mark.parametrize("shape", [(4, 32, 32, 64, 64), (8, 16, 16, 128, 128), (3, 16, 16, 64, 128), (64, 32, 32, 64, 64)])
def test_conv_wgrad_halo_kernel(ops, shape):
    """With a scratch buffer the 3x3 stride-1 wgrad on 16/32-pixel images takes the halo-reuse kernel."""
    n, h, w, cin, cout = shape
    x = nhwc_bf16(n, h, w, cin, 30)
    dy = nhwc_bf16(n, h, w, cout, 31, scale=0.2)
    ref = torch.nn.grad.conv2d_weight(to_nchw(x), (cout, cin, 3, 3), to_nchw(dy), padding=1)
    d = ops.conv_desc(n, h, w, cin, cout, 3, 1, flags=2)
    for floats in (ops.WGRAD_SCRATCH_FLOATS, 3 * (cin // 64) * (cout // 64) * 5 * 128 * 64):   # full and tight scratch
        scratch = torch.full((floats,), float("nan"), device="cuda")
        dw = torch.zeros(cout, cin, 3, 3, device="cuda")
        ops.conv_wgrad(d, x, dy, dw, scratch=scratch)
        torch.cuda.synchronize()
        assert rel_err(dw, ref) < F32_TOL
        ops.conv_wgrad(d, x, dy, dw, scratch=scratch)              # accumulates like the atomic kernel
        torch.cuda.synchronize()
        assert rel_err(dw, 2 * ref) < F32_TOL
    dw2 = torch.zeros(cout, cin, 3, 3, device="cuda")
    ops.conv_wgrad(d, x, dy, dw2)                                  # no scratch: atomic kernel, same answer
    torch.cuda.synchronize()
    assert rel_err(dw2, ref) < F32_TOL


def test_conv_wgrad_large_k_split(ops):
    n, h, w, cin, cout = 128, 32, 32, 64, 64
    x = nhwc_bf16(n, h, w, cin, 12)
    dy = nhwc_bf16(n, h, w, cout, 13, scale=0.1)
    ref = torch.nn.grad.conv2d_weight(to_nchw(x), (cout, cin, 3, 3), to_nchw(dy), padding=1)
    dw = torch.zeros(cout, cin, 3, 3, device="cuda")
    ops.conv_wgrad(ops.conv_desc(n, h, w, cin, cout, 3, 1), x, dy, dw)
    torch.cuda.synchronize()
    assert rel_err(dw, ref) < F32_TOL


def stem_operands(ops, n, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn(n, 3, 32, 32, device="cuda", generator=g)
    col = torch.empty(n, 32, 32, 64, device="cuda", dtype=torch.bfloat16)
    ops.stem_im2col(x, col)
    return x, col


def test_stem_im2col_and_conv(ops):
    n = 8
    x, col = stem_operands(ops, n, 14)
    torch.cuda.synchronize()
    xb = x.to(torch.bfloat16).float()
    unf = F.unfold(xb, 3, padding=1)                       # [n, c*9, L], index c*9 + tap
    unf = unf.view(n, 3, 9, 32, 32).permute(0, 3, 4, 2, 1).reshape(n, 32, 32, 27)   # k = tap*3 + c
    assert torch.equal(col[..., :27].float(), unf)
    assert (col[..., 27:] == 0).all()
    wt = weight(64, 3, 3, 15)
    wf = torch.zeros(64, 64, device="cuda", dtype=torch.bfloat16)
    wf[:, :27] = wt.permute(0, 2, 3, 1).reshape(64, 27).to(torch.bfloat16)
    y = torch.empty(n, 32, 32, 64, device="cuda", dtype=torch.bfloat16)
    ops.conv_fprop(ops.conv_desc(n, 32, 32, 64, 64, 1, 1), col, wf.view(1, 64, 64), y)
    torch.cuda.synchronize()
    ref = F.conv2d(xb, wt, padding=1)
    assert rel_err(to_nchw(y), ref) < BF16_TOL
    # stem wgrad through the flattened-k addressing
    dy = nhwc_bf16(n, 32, 32, 64, 16)
    dw = torch.zeros(64, 3, 3, 3, device="cuda")
    ops.conv_wgrad(ops.conv_desc(n, 32, 32, 64, 64, 1, 1, cin_valid=27, kflat_cin=3), col, dy, dw)
    torch.cuda.synchronize()
    refw = torch.nn.grad.conv2d_weight(xb, (64, 3, 3, 3), to_nchw(dy), padding=1)
    assert rel_err(dw, refw) < F32_TOL


@pytest.mark.parametrize("C,rows", [(64, 4096), (128, 1000), (512, 80)])
def test_bn_forward_chain(ops, C, rows):
    """partials -> finalize -> apply (+residual variants) against F.batch_norm."""
    g = torch.Generator(device="cuda").manual_seed(20)
    y = (torch.randn(rows, C, device="cuda", generator=g) * 2 + 0.5).to(torch.bfloat16)
    yf = y.float()
    nrows_p = 7
    chunks = torch.tensor_split(yf, nrows_p, 0)
    ps = torch.stack([c.sum(0) for c in chunks])
    pq = torch.stack([(c * c).sum(0) for c in chunks])
    gamma = torch.rand(C, device="cuda", generator=g) + 0.5
    beta = torch.randn(C, device="cuda", generator=g)
    rm = torch.randn(C, device="cuda", generator=g)
    rv = torch.rand(C, device="cuda", generator=g) + 0.5
    rm_ref, rv_ref = rm.clone(), rv.clone()
    scale, shift, mean, rstd = (torch.empty(C, device="cuda") for _ in range(4))
    ops.bn_finalize(ps, pq, nrows_p, C, rows, gamma, beta, 1e-5, 0.1, rm, rv, scale, shift, mean, rstd)
    ref = F.batch_norm(yf, rm_ref, rv_ref, gamma, beta, training=True, momentum=0.1, eps=1e-5)
    torch.cuda.synchronize()
    assert rel_err(rm, rm_ref) < 1e-5 and rel_err(rv, rv_ref) < 1e-5
    assert rel_err(mean, yf.mean(0)) < 1e-4
    out = torch.empty(rows, C, device="cuda", dtype=torch.bfloat16)
    ops.bn_act_apply(y, scale, shift, out, rows, C, relu=True)
    assert rel_err(out.float(), F.relu(ref)) < BF16_TOL
    res = (torch.randn(rows, C, device="cuda", generator=g)).to(torch.bfloat16)
    ops.bn_act_apply(y, scale, shift, out, rows, C, relu=True, res=res)
    assert rel_err(out.float(), F.relu(ref + res.float())) < BF16_TOL
    rs = torch.rand(C, device="cuda", generator=g) + 0.5
    rb = torch.randn(C, device="cuda", generator=g)
    ops.bn_act_apply(y, scale, shift, out, rows, C, relu=False, res=res, res_scale=rs, res_shift=rb)
    assert rel_err(out.float(), ref + res.float() * rs + rb) < BF16_TOL
    # eval-mode scale/shift
    ops.bn_eval_scale_shift(C, gamma, beta, rm, rv, 1e-5, scale, shift)
    ops.bn_act_apply(y, scale, shift, out, rows, C, relu=False)
    ref_eval = F.batch_norm(yf, rm, rv, gamma, beta, training=False, eps=1e-5)
    assert rel_err(out.float(), ref_eval) < BF16_TOL


@pytest.mark.parametrize("C,rows", [(64, 8192), (256, 1000), (512, 2048)])
def test_bn_relu_backward(ops, C, rows):
    g = torch.Generator(device="cuda").manual_seed(21)
    y = (torch.randn(rows, C, device="cuda", generator=g) * 1.5 + 0.3).to(torch.bfloat16)
    gamma = (torch.rand(C, device="cuda", generator=g) + 0.5).requires_grad_()
    beta = (torch.randn(C, device="cuda", generator=g) * 0.3).requires_grad_()
    dout = torch.randn(rows, C, device="cuda", generator=g).to(torch.bfloat16)
    yf = y.float().requires_grad_()
    act_ref = F.relu(F.batch_norm(yf, None, None, gamma, beta, training=True, eps=1e-5))
    act_ref.backward(dout.float())
    mean = y.float().mean(0)
    var = y.float().var(0, unbiased=False)
    rstd = (var + 1e-5).rsqrt()
    act = act_ref.detach().to(torch.bfloat16)
    prow = ops.bn_bwd_partial_rows(rows, C)
    partial = torch.empty(prow, 2, C, device="cuda")
    dz = torch.empty_like(dout)
    ops.bn_bwd_reduce(dout, act, y, dz, rows, C, partial)
    dgamma, dbeta, ca, cb, cc = (torch.empty(C, device="cuda") for _ in range(5))
    ops.bn_bwd_finalize(partial, prow, C, rows, gamma.detach(), mean, rstd, dgamma, dbeta, ca, cb, cc)
    dy = torch.empty_like(dout)
    ops.bn_bwd_apply(dz, y, ca, cb, cc, dy, rows, C)
    torch.cuda.synchronize()
    # the mask comes from the bf16-rounded activation: elements whose fp32 activation rounds to 0 differ; rare
    assert rel_err(dgamma, gamma.grad) < 5e-3
    assert rel_err(dbeta, beta.grad) < 5e-3
    assert rel_err(dy.float(), yf.grad) < BF16_TOL
    # in-place masking (dz_out aliases dout) and the no-activation form
    d2 = dout.clone()
    ops.bn_bwd_reduce(d2, act, y, d2, rows, C, partial)
    torch.cuda.synchronize()
    assert torch.equal(d2, dz)
    ops.bn_bwd_reduce(dz, None, y, None, rows, C, partial)
    ops.bn_bwd_finalize(partial, prow, C, rows, gamma.detach(), mean, rstd, dgamma, dbeta, ca, cb, cc)
    torch.cuda.synchronize()
    assert rel_err(dbeta, beta.grad) < 5e-3
    # mask recomputed from y (no act read, no dz written): the forward kernel's own act decides the expected mask,
    # so the result must equal the act-based path bit for bit
    scale = (gamma.detach() * rstd).contiguous()
    shift = (beta.detach() - mean * scale).contiguous()
    act_k = torch.empty_like(y)
    ops.bn_act_apply(y, scale, shift, act_k, rows, C, relu=True)
    ref = [torch.empty(C, device="cuda") for _ in range(5)]
    new = [torch.empty(C, device="cuda") for _ in range(5)]
    dy_ref, dy_new = torch.empty_like(dout), torch.empty_like(dout)
    ops.bn_bwd_reduce(dout, act_k, y, dz, rows, C, partial)
    ops.bn_bwd_finalize(partial, prow, C, rows, gamma.detach(), mean, rstd, *ref)
    ops.bn_bwd_apply(dz, y, ref[2], ref[3], ref[4], dy_ref, rows, C)
    ops.bn_bwd_reduce(dout, None, y, None, rows, C, partial, scale, shift)
    ops.bn_bwd_finalize(partial, prow, C, rows, gamma.detach(), mean, rstd, *new)
    ops.bn_bwd_apply(dout, y, new[2], new[3], new[4], dy_new, rows, C, relu_scale=scale, relu_shift=shift)
    torch.cuda.synchronize()
    for a, b in zip(ref, new):
        assert torch.equal(a, b)
    assert torch.equal(dy_ref, dy_new)
    # mask taken from the bit mask the forward apply pass writes (residual units): same sums, same dz
    bits = torch.zeros(rows * C // 8, dtype=torch.uint8, device="cuda")
    res = torch.randn(rows, C, device="cuda", generator=g).to(torch.bfloat16)
    ops.bn_act_apply(y, scale, shift, act_k, rows, C, relu=True, res=res, relu_bits=bits)
    w8 = (1 << torch.arange(8, device="cuda")).to(torch.int32)
    assert torch.equal(bits, ((act_k.view(-1, 8) > 0).to(torch.int32) * w8).sum(1).to(torch.uint8))
    dz_a, dz_b = torch.empty_like(dout), torch.empty_like(dout)
    pa, pb = torch.empty_like(partial), torch.empty_like(partial)
    ops.bn_bwd_reduce(dout, act_k, y, dz_a, rows, C, pa)
    ops.bn_bwd_reduce(dout, None, y, dz_b, rows, C, pb, relu_bits=bits)
    torch.cuda.synchronize()
    assert torch.equal(dz_a, dz_b) and torch.equal(pa, pb)


def test_head_and_loss(ops):
    n, hw, C, K = 128, 16, 512, 10
    g = torch.Generator(device="cuda").manual_seed(22)
    act = torch.rand(n, hw, C, device="cuda", generator=g).to(torch.bfloat16)
    W = (torch.randn(K, C, device="cuda", generator=g) * 0.05).requires_grad_()
    b = (torch.randn(K, device="cuda", generator=g) * 0.1).requires_grad_()
    target = torch.randint(0, K, (n,), device="cuda", generator=g)
    pooled = torch.empty(n, C, device="cuda")
    logits = torch.empty(n, K, device="cuda")
    ops.head_fwd(act, n, hw, C, W.detach(), b.detach(), K, pooled, logits)
    af = act.float().requires_grad_()
    ref_logits = F.linear(af.mean(1), W, b)
    loss_ref = F.cross_entropy(ref_logits, target)
    loss_ref.backward()
    dlogits = torch.empty(n, K, device="cuda")
    loss = torch.empty(1, device="cuda")
    correct = torch.zeros(1, device="cuda", dtype=torch.int32)
    ops.ce_loss(logits, target, n, K, dlogits, loss, correct)
    dact = torch.empty(n, hw, C, device="cuda", dtype=torch.bfloat16)
    dW = torch.empty(K, C, device="cuda")
    db = torch.empty(K, device="cuda")
    ops.head_bwd(dlogits, pooled, W.detach(), n, hw, C, K, dact, dW, db)
    torch.cuda.synchronize()
    assert rel_err(logits, ref_logits) < 1e-4
    assert abs(loss.item() - loss_ref.item()) < 1e-4
    assert correct.item() == (ref_logits.argmax(1) == target).sum().item()
    assert rel_err(dW, W.grad) < 1e-4 and rel_err(db, b.grad) < 1e-4
    assert rel_err(dact.float(), af.grad) < BF16_TOL


def test_sgd_step(ops):
    nparam = 4096 * 3 + 8
    g = torch.Generator(device="cuda").manual_seed(23)
    p = torch.randn(nparam, device="cuda", generator=g)
    ref_p = p.clone().requires_grad_()
    opt = torch.optim.SGD([ref_p], lr=0.1, momentum=0.9, weight_decay=5e-4)
    buf = torch.zeros_like(p)
    hyper = torch.tensor([0.1, 0.9, 5e-4, 1.0, 1.0], device="cuda")
    for step in range(3):
        grad = torch.randn(nparam, device="cuda", generator=g)
        ref_p.grad = grad.clone()
        opt.step()
        hyper[4] = 1.0 if step == 0 else 0.0
        ops.sgd_step(p, grad, buf, hyper)
    torch.cuda.synchronize()
    assert (p - ref_p.detach()).abs().max().item() < 1e-6


def test_repack_weights(ops):
    g = torch.Generator(device="cuda").manual_seed(24)
    w1 = torch.randn(128, 64, 3, 3, device="cuda", generator=g)
    w2 = torch.randn(128, 64, 1, 1, device="cuda", generator=g)
    w0 = torch.randn(64, 3, 3, 3, device="cuda", generator=g)
    master = torch.cat([w1.flatten(), w2.flatten(), w0.flatten()])
    wf = torch.zeros(w1.numel() + w2.numel() + 64 * 64, device="cuda", dtype=torch.bfloat16)
    wd = torch.zeros(w1.numel() + w2.numel(), device="cuda", dtype=torch.bfloat16)
    entries = [
        dict(src_off=0, wf_off=0, wd_off=0, cout=128, cin=64, taps=9, cin_pad=64),
        dict(src_off=w1.numel(), wf_off=w1.numel(), wd_off=w1.numel(), cout=128, cin=64, taps=1, cin_pad=64),
        dict(src_off=w1.numel() + w2.numel(), wf_off=w1.numel() + w2.numel(), wd_off=-1, cout=64, cin=3, taps=9,
             cin_pad=64, kflat=1),
    ]
    table = ops.make_repack_table(entries, "cuda")
    ops.repack_weights(master, wf, wd, table, len(entries))
    torch.cuda.synchronize()
    assert torch.equal(wf[:w1.numel()].view(9, 128, 64), ops.pack_weight_fprop(w1))
    assert torch.equal(wd[:w1.numel()].view(9, 64, 128), ops.pack_weight_dgrad(w1))
    assert torch.equal(wf[w1.numel():w1.numel() + w2.numel()].view(1, 128, 64), ops.pack_weight_fprop(w2))
    stem = wf[w1.numel() + w2.numel():].view(64, 64)
    assert torch.equal(stem[:, :27], w0.permute(0, 2, 3, 1).reshape(64, 27).to(torch.bfloat16))
    assert (stem[:, 27:] == 0).all()


# ------------------------------------------------------------------------------------------------
# Guard bands: every output of the newer kernels sits inside a larger sentinel-filled allocation; nothing outside
# the tensor may change (compute-sanitizer is not available on the GPU pool, so out-of-bounds stores are caught here).
# ------------------------------------------------------------------------------------------------
class _Guarded:
    PAD = 4096        # elements on either side

    def __init__(self, shape, dtype, fill):
        n = int(np.prod(shape))
        self.raw = torch.empty(n + 2 * self.PAD, device="cuda", dtype=dtype)
        self.sentinel = 123 if dtype == torch.uint8 else -7777.0
        self.raw.fill_(self.sentinel)
        self.t = self.raw[self.PAD:self.PAD + n].view(shape)
        self.t.fill_(fill)

    def intact(self):
        return bool((self.raw[:self.PAD] == self.sentinel).all() and (self.raw[-self.PAD:] == self.sentinel).all())


def test_outputs_stay_inside_their_tensors(ops):
    n, h, cin, cout = 5, 16, 128, 256                 # odd batch: ragged last tile in every kernel
    g = torch.Generator(device="cuda").manual_seed(71)
    bad = []

    def check(name, *gs):
        torch.cuda.synchronize()
        for i, gd in enumerate(gs):
            if not gd.intact():
                bad.append(f"{name}[{i}]")

    # fprop (+stats), dgrad (+add), fused dgrad+BN, stride-2 dgrad with shortcut
    for (hh, ci, co, k, s) in [(16, 128, 128, 3, 1), (8, 256, 256, 3, 1), (32, 64, 64, 3, 1), (16, 128, 256, 3, 2),
                               (16, 128, 256, 1, 2), (4, 512, 512, 3, 1), (32, 64, 64, 1, 1)]:
        ho = hh // s
        x = nhwc_bf16(n, hh, hh, ci, 72)
        wt = weight(co, ci, k, 73)
        d = ops.conv_desc(n, hh, hh, ci, co, k, s)
        y = _Guarded((n, ho, ho, co), torch.bfloat16, 0.0)
        ps, pq = _Guarded((148, co), torch.float32, 0.0), _Guarded((148, co), torch.float32, 0.0)
        ops.conv_fprop(d, x, ops.pack_weight_fprop(wt), y.t, ps.t, pq.t)
        check(f"fprop{(hh, ci, co, k, s)}", y, ps, pq)
        dy = nhwc_bf16(n, ho, ho, co, 74)
        dx = _Guarded((n, hh, hh, ci), torch.bfloat16, 0.0)
        ops.conv_dgrad(d, dy, ops.pack_weight_dgrad(wt), dx.t)
        check(f"dgrad{(hh, ci, co, k, s)}", dx)
        if not (k == 1 and s == 2):
            dz = _Guarded((n, hh, hh, ci), torch.bfloat16, 0.0)
            part = _Guarded((148, 2, ci), torch.float32, 0.0)
            sc = torch.rand(ci, device="cuda", generator=g) + 0.5
            sh = torch.randn(ci, device="cuda", generator=g)
            ops.conv_dgrad_bn(d, dy, ops.pack_weight_dgrad(wt), dz.t, x, sc, sh, part.t)
            check(f"dgrad_bn{(hh, ci, co, k, s)}", dz, part)
        if k == 3 and s == 2:
            w10 = torch.cat([ops.pack_weight_dgrad(wt), ops.pack_weight_dgrad(weight(co, ci, 1, 75))], 0).contiguous()
            dx2 = _Guarded((n, hh, hh, ci), torch.bfloat16, 0.0)
            ops.conv_dgrad_with_shortcut(d, dy, nhwc_bf16(n, ho, ho, co, 76), w10, dx2.t)
            check("dgrad_shortcut", dx2)
        for pf in (128, 129):                          # CTA-pair kernels (halo / per-tap)
            dp = ops.conv_desc(n, hh, hh, ci, co, k, s, flags=pf)
            y2 = _Guarded((n, ho, ho, co), torch.bfloat16, 0.0)
            ops.conv_fprop(dp, x, ops.pack_weight_fprop(wt), y2.t, ps.t, pq.t)
            check(f"pair_fprop{(hh, ci, co, k, s, pf)}", y2, ps, pq)
            if s == 1:
                dx3 = _Guarded((n, hh, hh, ci), torch.bfloat16, 0.0)
                ops.conv_dgrad(dp, dy, ops.pack_weight_dgrad(wt), dx3.t)
                dz3 = _Guarded((n, hh, hh, ci), torch.bfloat16, 0.0)
                part3 = _Guarded((148, 2, ci), torch.float32, 0.0)
                ops.conv_dgrad_bn(dp, dy, ops.pack_weight_dgrad(wt), dz3.t, x, sc, sh, part3.t)
                check(f"pair_dgrad{(hh, ci, co, k, s, pf)}", dx3, dz3, part3)
        for flags in (0, 32, 32 | 64, 32 | 128 | 256):
            dw = _Guarded((co * ci * k * k,), torch.float32, 0.0)
            ops.conv_wgrad(ops.conv_desc(n, hh, hh, ci, co, k, s, flags=flags), x, dy, dw.t)
            check(f"wgrad{(hh, ci, co, k, s, flags)}", dw)
    # BN apply with bit mask, BN backward reduce/apply
    rows, C = n * 16 * 16, 128
    yb = nhwc_bf16(n, 16, 16, C, 77).view(rows, C)
    sc = torch.rand(C, device="cuda", generator=g) + 0.5
    sh = torch.randn(C, device="cuda", generator=g)
    out, bits = _Guarded((rows, C), torch.bfloat16, 0.0), _Guarded((rows * C // 8,), torch.uint8, 0)
    ops.bn_act_apply(yb, sc, sh, out.t, rows, C, relu=True, res=yb, relu_bits=bits.t)
    check("bn_act_apply", out, bits)
    prow = ops.bn_bwd_partial_rows(rows, C)
    part, dzg = _Guarded((prow, 2, C), torch.float32, 0.0), _Guarded((rows, C), torch.bfloat16, 0.0)
    ops.bn_bwd_reduce(yb, None, yb, dzg.t, rows, C, part.t, relu_bits=bits.t)
    check("bn_bwd_reduce", part, dzg)
    # depthwise strip kernels
    for (hh, C) in [(32, 96), (16, 192), (8, 384), (4, 960)]:
        x = nhwc_bf16(n, hh, hh, C, 78)
        wt = torch.randn(C, 1, 3, 3, device="cuda", generator=g)
        y = _Guarded((n, hh, hh, C), torch.bfloat16, 0.0)
        ps, pq = _Guarded((ops.MAX_DW_STAT_ROWS, C), torch.float32, 0.0), _Guarded((ops.MAX_DW_STAT_ROWS, C), torch.float32, 0.0)
        ops.dwconv_fprop(x, wt, y.t, C, 1, ps.t, pq.t)
        dx = _Guarded((n, hh, hh, C), torch.bfloat16, 0.0)
        ops.dwconv_dgrad(x, wt, dx.t, C, 1)
        dw = _Guarded((C, 1, 3, 3), torch.float32, 0.0)
        ops.dwconv_wgrad(x, x, dw.t, C, 1)
        check(f"dw{(hh, C)}", y, ps, pq, dx, dw)
        if hh >= 8:                                   # stride-2 strip kernels
            y2 = _Guarded((n, hh // 2, hh // 2, C), torch.bfloat16, 0.0)
            ops.dwconv_fprop(x, wt, y2.t, C, 2, ps.t, pq.t)
            dx2 = _Guarded((n, hh, hh, C), torch.bfloat16, 0.0)
            ops.dwconv_dgrad(y2.t, wt, dx2.t, C, 2)
            dw2 = _Guarded((C, 1, 3, 3), torch.float32, 0.0)
            ops.dwconv_wgrad(x, y2.t, dw2.t, C, 2)
            check(f"dw_s2{(hh, C)}", y2, ps, pq, dx2, dw2)
    assert not bad, bad


@pytest.mark.parametrize("shape", [(32, 4, 4, 512, 512, 3, 1), (8, 16, 16, 128, 128, 3, 1), (4, 32, 32, 64, 64, 3, 1),
                                   (2, 32, 32, 64, 64, 1, 1), (5, 4, 4, 512, 512, 3, 1), (3, 8, 8, 256, 256, 3, 1),
                                   (64, 8, 8, 128, 256, 3, 2), (300, 4, 4, 256, 256, 3, 1)])
def test_conv_cta_pair_kernel(ops, shape):
    """flags bit 7: the CTA-pair kernels (tcgen05.mma.cta_group::2) against their single-CTA counterparts on the same
    operands — conv_halo2.cu vs conv_halo.cu (flags 128 vs 0), conv_gemm2.cu vs conv_gemm.cu (flags 129 vs 1): same K
    order and fp32 accumulation, so outputs and BatchNorm sums agree to the last bit's rounding; odd tile counts
    leave the peer CTA of the last pair without a tile."""
    n, h, w, cin, cout, k, stride = shape
    ho, wo = h // stride, w // stride
    x = nhwc_bf16(n, h, w, cin, 81)
    wt = weight(cout, cin, k, 82)
    wf, wd = ops.pack_weight_fprop(wt), ops.pack_weight_dgrad(wt)
    ref = F.conv2d(to_nchw(x), wt.to(torch.bfloat16).float(), stride=stride, padding=k // 2)
    dy = nhwc_bf16(n, ho, wo, cout, 83)
    add = nhwc_bf16(n, h, w, cin, 84)
    for single, pair in ((1, 129), (0, 128)):
        outs, sums = [], []
        for flags in (single, pair):
            y = torch.full((n, ho, wo, cout), float("nan"), device="cuda", dtype=torch.bfloat16)
            ps, pq = torch.zeros(148, cout, device="cuda"), torch.zeros(148, cout, device="cuda")
            rows = ops.conv_fprop(ops.conv_desc(n, h, w, cin, cout, k, stride, flags=flags), x, wf, y, ps, pq)
            torch.cuda.synchronize()
            assert 1 <= rows <= 148
            outs.append(y); sums.append((ps[:rows].double().sum(0), pq[:rows].double().sum(0)))
        assert torch.isfinite(outs[1].float()).all(), pair
        assert rel_err(outs[1].float(), outs[0].float()) < 1e-5, pair
        assert rel_err(sums[1][0], sums[0][0]) < 1e-5 and rel_err(sums[1][1], sums[0][1]) < 1e-5, pair
        assert rel_err(to_nchw(outs[1]), ref) < BF16_TOL
        if stride == 1:
            dxs = []
            for flags in (single, pair):
                dx = torch.full((n, h, w, cin), float("nan"), device="cuda", dtype=torch.bfloat16)
                ops.conv_dgrad(ops.conv_desc(n, h, w, cin, cout, k, stride, flags=flags), dy, wd, dx, add)
                torch.cuda.synchronize()
                dxs.append(dx)
            assert torch.isfinite(dxs[1].float()).all(), pair
            assert rel_err(dxs[1].float(), dxs[0].float()) < 1e-5, pair


# ------------------------------------------------------------------------------------------------
# Statistics accumulated into one row + finalize folded into the apply passes (B2cConvDesc.flags bit 9,
# b2c_bn_act_apply_stats, b2c_bn_bwd_reduce_acc, b2c_bn_bwd_apply_stats) against the two / three-call forms.
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("shape", [(4, 32, 32, 64, 64, 3, 1), (64, 32, 32, 64, 64, 3, 1), (8, 16, 16, 128, 128, 3, 1),
                                   (16, 8, 8, 256, 256, 3, 1), (32, 4, 4, 512, 512, 3, 1), (8, 32, 32, 64, 128, 3, 2),
                                   (8, 32, 32, 64, 128, 1, 2), (5, 4, 4, 512, 512, 3, 1)])
@pytest.mark.parametrize("pair", [0, 128])
def test_conv_statistics_accumulated(ops, shape, pair):
    n, h, w, cin, cout, k, stride = shape
    x = nhwc_bf16(n, h, w, cin, 81)
    wf = ops.pack_weight_fprop(weight(cout, cin, k, 82))
    ho, wo = h // stride, w // stride
    y0 = torch.empty(n, ho, wo, cout, device="cuda", dtype=torch.bfloat16)
    ps = torch.zeros(ops.MAX_STAT_ROWS, cout, device="cuda")
    pq = torch.zeros(ops.MAX_STAT_ROWS, cout, device="cuda")
    rows = ops.conv_fprop(ops.conv_desc(n, h, w, cin, cout, k, stride, flags=pair), x, wf, y0, ps, pq)
    y1 = torch.empty_like(y0)
    acc = torch.zeros(2, cout, device="cuda")
    for rep in (1, 2):                                    # accumulates: a second call doubles the sums
        r1 = ops.conv_fprop(ops.conv_desc(n, h, w, cin, cout, k, stride, flags=pair | 512), x, wf, y1, acc[0], acc[1])
        torch.cuda.synchronize()
        assert r1 == 1 and torch.equal(y0, y1)
        assert (acc[0] - rep * ps[:rows].sum(0)).abs().max() <= 1e-4 * ps[:rows].abs().sum(0).max() + 1e-3
        assert rel_err(acc[1], rep * pq[:rows].sum(0)) < 1e-5
    # the fused dgrad + BatchNorm-backward reduction, same flag
    dy = nhwc_bf16(n, ho, wo, cout, 83)
    wd = ops.pack_weight_dgrad(weight(cout, cin, k, 82))
    if not (k == 1 and stride == 2):
        y_bn = nhwc_bf16(n, h, w, cin, 84)
        g = torch.Generator(device="cuda").manual_seed(85)
        scale = torch.rand(cin, device="cuda", generator=g) + 0.5
        shift = torch.randn(cin, device="cuda", generator=g) * 0.3
        dz0, dz1 = torch.empty(n, h, w, cin, device="cuda", dtype=torch.bfloat16), None
        part = torch.zeros(148, 2, cin, device="cuda")
        prow = ops.conv_dgrad_bn(ops.conv_desc(n, h, w, cin, cout, k, stride, flags=pair), dy, wd, dz0, y_bn, scale,
                                 shift, part)
        dz1 = torch.empty_like(dz0)
        sums = torch.zeros(2, cin, device="cuda")
        p1 = ops.conv_dgrad_bn(ops.conv_desc(n, h, w, cin, cout, k, stride, flags=pair | 512), dy, wd, dz1, y_bn,
                               scale, shift, sums)
        torch.cuda.synchronize()
        assert p1 == 1 and torch.equal(dz0, dz1)
        ref = part[:prow].sum(0)
        assert (sums - ref).abs().max() <= 1e-4 * part[:prow].abs().sum(0).max() + 1e-3


@pytest.mark.parametrize("C,rows", [(64, 4096), (192, 1000), (512, 80), (1280, 333), (64, 131072)])
def test_bn_apply_with_fused_finalize(ops, C, rows):
    """b2c_bn_act_apply_stats == b2c_bn_finalize + b2c_bn_act_apply: same scale / shift / mean / rstd / running
    statistics and the same output bits, with identity and BatchNorm'ed residuals and the ReLU bit mask."""
    g = torch.Generator(device="cuda").manual_seed(90)
    y = (torch.randn(rows, C, device="cuda", generator=g) * 2 + 0.5).to(torch.bfloat16)
    r = torch.randn(rows, C, device="cuda", generator=g).to(torch.bfloat16)

    def bn_inputs(t, seed):
        gg = torch.Generator(device="cuda").manual_seed(seed)
        tf = t.float()
        d = dict(sum=tf.sum(0).contiguous(), sq=(tf * tf).sum(0).contiguous(),
                 gamma=torch.rand(C, device="cuda", generator=gg) + 0.5, beta=torch.randn(C, device="cuda", generator=gg),
                 rm=torch.randn(C, device="cuda", generator=gg), rv=torch.rand(C, device="cuda", generator=gg) + 0.5)
        return d

    def two_call(d):
        o = {k: torch.empty(C, device="cuda") for k in ("scale", "shift", "mean", "rstd")}
        rm, rv = d["rm"].clone(), d["rv"].clone()
        ops.bn_finalize(d["sum"].view(1, C), d["sq"].view(1, C), 1, C, rows, d["gamma"], d["beta"], 1e-5, 0.1, rm, rv,
                        o["scale"], o["shift"], o["mean"], o["rstd"])
        return o, rm, rv

    def record(d):
        o = {k: torch.empty(C, device="cuda") for k in ("scale", "shift", "mean", "rstd")}
        rm, rv = d["rm"].clone(), d["rv"].clone()
        rec = ops.bn_stats(d["sum"], d["sq"], d["gamma"], d["beta"], rows, 1e-5, 0.1, rm, rv, o["scale"], o["shift"],
                           o["mean"], o["rstd"])
        return rec, o, rm, rv
    dm, dr = bn_inputs(y, 91), bn_inputs(r, 92)
    (om, rm0, rv0), (orr, _, _) = two_call(dm), two_call(dr)
    for variant in ("plain", "identity", "affine"):
        out0, out1 = torch.empty_like(y), torch.empty_like(y)
        bits0 = torch.zeros(rows * C // 8, dtype=torch.uint8, device="cuda")
        bits1 = torch.zeros_like(bits0)
        rec, o1, rm1, rv1 = record(dm)
        if variant == "plain":
            ops.bn_act_apply(y, om["scale"], om["shift"], out0, rows, C, relu=True)
            ops.bn_act_apply_stats(y, rec, out1, rows, C, relu=True)
        elif variant == "identity":
            ops.bn_act_apply(y, om["scale"], om["shift"], out0, rows, C, relu=True, res=r, relu_bits=bits0)
            ops.bn_act_apply_stats(y, rec, out1, rows, C, relu=True, res=r, relu_bits=bits1)
        else:
            recr, o2, _, _ = record(dr)
            ops.bn_act_apply(y, om["scale"], om["shift"], out0, rows, C, relu=False, res=r, res_scale=orr["scale"],
                             res_shift=orr["shift"])
            ops.bn_act_apply_stats(y, rec, out1, rows, C, relu=False, res=r, res_bn=recr)
            torch.cuda.synchronize()
            assert torch.equal(o2["scale"], orr["scale"]) and torch.equal(o2["shift"], orr["shift"])
        torch.cuda.synchronize()
        for kname in om:
            assert torch.equal(o1[kname], om[kname]), (variant, kname)
        assert torch.equal(rm1, rm0) and torch.equal(rv1, rv0)
        assert torch.equal(out0, out1) and torch.equal(bits0, bits1), variant


@pytest.mark.parametrize("C,rows", [(64, 4096), (192, 1000), (512, 80), (1280, 333)])
def test_bn_backward_with_fused_finalize(ops, C, rows):
    """b2c_bn_bwd_reduce_acc + b2c_bn_bwd_apply_stats == b2c_bn_bwd_reduce + b2c_bn_bwd_finalize + b2c_bn_bwd_apply."""
    g = torch.Generator(device="cuda").manual_seed(95)
    y = (torch.randn(rows, C, device="cuda", generator=g) * 1.5 + 0.3).to(torch.bfloat16)
    dout = torch.randn(rows, C, device="cuda", generator=g).to(torch.bfloat16)
    add = torch.randn(rows, C, device="cuda", generator=g).to(torch.bfloat16)
    gamma = torch.rand(C, device="cuda", generator=g) + 0.5
    beta = torch.randn(C, device="cuda", generator=g) * 0.3
    mean = y.float().mean(0)
    rstd = (y.float().var(0, unbiased=False) + 1e-5).rsqrt()
    scale = (gamma * rstd).contiguous()
    shift = (beta - mean * scale).contiguous()
    prow = ops.bn_bwd_partial_rows(rows, C)
    partial = torch.empty(prow, 2, C, device="cuda")
    ref = [torch.empty(C, device="cuda") for _ in range(5)]
    dy0, dy1 = torch.empty_like(dout), torch.empty_like(dout)
    for masked in (False, True):
        sc, sh = (scale, shift) if masked else (None, None)
        ops.bn_bwd_reduce(dout, None, y, None, rows, C, partial, sc, sh)
        ops.bn_bwd_finalize(partial, prow, C, rows, gamma, mean, rstd, *ref)
        ops.bn_bwd_apply(dout, y, ref[2], ref[3], ref[4], dy0, rows, C, add=add, relu_scale=sc, relu_shift=sh)
        sums = torch.zeros(2, C, device="cuda")
        dg, db = torch.empty(C, device="cuda"), torch.empty(C, device="cuda")
        ops.bn_bwd_reduce_acc(dout, None, y, None, rows, C, sums, sc, sh)
        ops.bn_bwd_apply_stats(dout, y, sums, rows, gamma, mean, rstd, dg, db, dy1, rows, C, add=add, relu_scale=sc,
                               relu_shift=sh)
        torch.cuda.synchronize()
        want = partial.double().sum(0)
        assert (sums.double() - want).abs().max() <= 1e-5 * partial.abs().sum(0).max() + 1e-4
        assert rel_err(dg, ref[0]) < 1e-4 and rel_err(db, ref[1]) < 1e-4
        assert rel_err(dy1.float(), dy0.float()) < 2e-3           # coefficients from sums that differ by round-off


@pytest.mark.parametrize("shape", [(8, 32, 32, 64, 64), (16, 16, 16, 128, 128), (32, 8, 8, 256, 256), (64, 4, 4, 512, 512),
                                   (5, 4, 4, 512, 512)])
def test_conv1x1_over_concat(ops, shape):
    """b2c_conv1x1_cat_fprop: the Root of a DLA tree, conv1x1(cat([x0, x1], 1)) (models/dla_simple.py:52-53) read
    through two tensor maps, against F.conv2d on the materialised concatenation; its gradients are the ordinary 1x1
    dgrad / wgrad with one weight slab each."""
    n, h, w, c, cout = shape
    x0, x1 = nhwc_bf16(n, h, w, c, 101), nhwc_bf16(n, h, w, c, 102)
    wt = weight(cout, 2 * c, 1, 103)
    ref = F.conv2d(torch.cat([to_nchw(x0), to_nchw(x1)], 1), wt)
    w2 = torch.stack([wt[:, :c, 0, 0], wt[:, c:, 0, 0]]).contiguous().to(torch.bfloat16)        # [2][cout][c]
    y = torch.full((n, h, w, cout), float("nan"), device="cuda", dtype=torch.bfloat16)
    ps = torch.zeros(ops.MAX_STAT_ROWS, cout, device="cuda")
    pq = torch.zeros(ops.MAX_STAT_ROWS, cout, device="cuda")
    rows = ops.conv1x1_cat_fprop(ops.conv_desc(n, h, w, c, cout, 1, 1), x0, x1, w2, y, ps, pq)
    torch.cuda.synchronize()
    assert torch.isfinite(to_nchw(y)).all() and rel_err(to_nchw(y), ref) < BF16_TOL
    assert rel_err(pq[:rows].sum(0), (ref * ref).sum((0, 2, 3))) < F32_TOL
    acc = torch.zeros(2, cout, device="cuda")
    y2 = torch.empty_like(y)
    assert ops.conv1x1_cat_fprop(ops.conv_desc(n, h, w, c, cout, 1, 1, flags=512), x0, x1, w2, y2, acc[0], acc[1]) == 1
    torch.cuda.synchronize()
    assert torch.equal(y, y2) and rel_err(acc[1], pq[:rows].sum(0)) < 1e-5
    # gradients slab by slab
    dy = nhwc_bf16(n, h, w, cout, 104)
    gx = torch.nn.grad.conv2d_input((n, 2 * c, h, w), wt, to_nchw(dy))
    gw = torch.nn.grad.conv2d_weight(torch.cat([to_nchw(x0), to_nchw(x1)], 1), (cout, 2 * c, 1, 1), to_nchw(dy))
    for t, xt in enumerate((x0, x1)):
        wd = wt[:, t * c:(t + 1) * c, 0, 0].t().contiguous().to(torch.bfloat16).view(1, c, cout)
        dx = torch.empty(n, h, w, c, device="cuda", dtype=torch.bfloat16)
        ops.conv_dgrad(ops.conv_desc(n, h, w, c, cout, 1, 1), dy, wd, dx)
        dw = torch.zeros(1, c, cout, device="cuda")
        ops.conv_wgrad(ops.conv_desc(n, h, w, c, cout, 1, 1, flags=32), xt, dy, dw)
        torch.cuda.synchronize()
        assert rel_err(to_nchw(dx), gx[:, t * c:(t + 1) * c]) < BF16_TOL
        assert rel_err(dw[0].t(), gw[:, t * c:(t + 1) * c, 0, 0]) < 2e-3


@pytest.mark.parametrize("shape", [(6, 16, 16, 128, 128, 4), (5, 8, 8, 256, 256, 4), (3, 16, 16, 128, 128, 3),
                                   (130, 8, 8, 256, 256, 4)])
def test_conv1x1_over_a_concat_of_up_to_four(ops, shape):
    """b2c_conv1x1_catn_fprop: the Root of a level-2 tree of models/dla.py:61,72-84 — conv1x1(cat(xs, 1)) over FOUR
    tensors, each read through its own tensor map — against F.conv2d on the materialised concatenation; statistics
    like b2c_conv_fprop; with two sources it is b2c_conv1x1_cat_fprop bit for bit."""
    n, h, w, c, cout, ns = shape
    xs = [nhwc_bf16(n, h, w, c, 140 + t) for t in range(ns)]
    wt = weight(cout, ns * c, 1, 150)
    ref = F.conv2d(torch.cat([to_nchw(x) for x in xs], 1), wt)
    wn = torch.stack([wt[:, t * c:(t + 1) * c, 0, 0] for t in range(ns)]).contiguous().to(torch.bfloat16)    # [ns][cout][c]
    whole, y, pad = _guarded((n, h, w, cout), torch.bfloat16, 555.0)
    ps = torch.zeros(ops.MAX_STAT_ROWS, cout, device="cuda")
    pq = torch.zeros(ops.MAX_STAT_ROWS, cout, device="cuda")
    rows = ops.conv1x1_catn_fprop(ops.conv_desc(n, h, w, c, cout, 1, 1), xs, wn, y, ps, pq)
    torch.cuda.synchronize()
    assert _guards_intact(whole, pad, 555.0)
    assert torch.isfinite(to_nchw(y)).all() and rel_err(to_nchw(y), ref) < BF16_TOL
    assert rel_err(ps[:rows].sum(0), ref.sum((0, 2, 3))) < 5e-3 or (ps[:rows].sum(0) - ref.sum((0, 2, 3))).abs().max() < 0.5
    assert rel_err(pq[:rows].sum(0), (ref * ref).sum((0, 2, 3))) < F32_TOL
    y2 = torch.empty_like(y)
    ops.conv1x1_catn_fprop(ops.conv_desc(n, h, w, c, cout, 1, 1), xs[:2], wn[:2].contiguous(), y2)
    y3 = torch.empty_like(y)
    ops.conv1x1_cat_fprop(ops.conv_desc(n, h, w, c, cout, 1, 1), xs[0], xs[1], wn[:2].contiguous(), y3)
    torch.cuda.synchronize()
    assert torch.equal(y2, y3)


@pytest.mark.parametrize("shape", [(8, 32, 32, 64, 128, 3, 2), (16, 16, 16, 128, 256, 3, 2), (64, 8, 8, 256, 512, 3, 2),
                                   (8, 32, 32, 64, 128, 1, 2), (5, 8, 8, 128, 128, 3, 2), (1024, 32, 32, 64, 128, 3, 2)])
def test_stride2_dgrad_on_cta_pairs(ops, shape):
    """flags bit 7 for the stride-2 data gradients: four output parity phases (3x3) / one strided phase (1x1) on the
    CTA-pair kernel; same result as the single-CTA kernel (same operands, same accumulation order)."""
    n, h, w, cin, cout, k, stride = shape
    dy = nhwc_bf16(n, h // 2, w // 2, cout, 111)
    wd = ops.pack_weight_dgrad(weight(cout, cin, k, 112))
    outs = []
    for flags in (0, 128):
        dx = torch.full((n, h, w, cin), float("nan"), device="cuda", dtype=torch.bfloat16)
        ops.conv_dgrad(ops.conv_desc(n, h, w, cin, cout, k, stride, flags=flags), dy, wd, dx)
        outs.append(dx)
    torch.cuda.synchronize()
    assert torch.isfinite(outs[1].float()).all() and rel_err(outs[1].float(), outs[0].float()) < 1e-3
    ref = torch.nn.grad.conv2d_input((n, cin, h, w), weight(cout, cin, k, 112), to_nchw(dy), stride=2, padding=k // 2)
    assert rel_err(to_nchw(outs[1]), ref) < BF16_TOL


def _guarded(shape, dtype, fill):
    """A tensor of `shape` carved out of the middle of a larger sentinel-filled allocation (256-byte aligned), so that a
    kernel writing outside its output is caught without a memory checker."""
    n = 1
    for s in shape:
        n *= s
    esz = torch.tensor([], dtype=dtype).element_size()
    pad = 4096 // esz
    whole = torch.full((n + 2 * pad,), fill, device="cuda", dtype=dtype)
    return whole, whole[pad:pad + n].view(shape), pad


def _guards_intact(whole, pad, fill):
    return bool((whole[:pad] == fill).all() and (whole[-pad:] == fill).all())


def test_round2_kernels_write_only_their_outputs(ops):
    """Guard bands around the outputs of the kernels added or changed in round 2 (ragged sizes on purpose): the
    CTA-pair stride-2 data gradients, the prefetching epilogue of the 64-wide halo kernel, the two-map 1x1 conv, the
    weight re-layout and the BatchNorm apply with fused finalize."""
    S = 777.0
    # stride-2 dgrad phases + fused shortcut on CTA pairs, odd tile count
    n, h, cin, cout = 5, 16, 64, 128
    dy, dys = nhwc_bf16(n, h // 2, h // 2, cout, 121), nhwc_bf16(n, h // 2, h // 2, cout, 122)
    w3, w1 = weight(cout, cin, 3, 123), weight(cout, cin, 1, 124)
    w10 = torch.cat([ops.pack_weight_dgrad(w3), ops.pack_weight_dgrad(w1)], 0).contiguous()
    whole, dx, pad = _guarded((n, h, h, cin), torch.bfloat16, S)
    ops.conv_dgrad_with_shortcut(ops.conv_desc(n, h, h, cin, cout, 3, 2, flags=128), dy, dys, w10, dx)
    torch.cuda.synchronize()
    assert _guards_intact(whole, pad, S) and torch.isfinite(dx.float()).all() and (dx != S).float().mean() > 0.99
    # 64-wide halo kernel, dgrad with a residual (epilogue operand requested a tile ahead), odd tile count
    n, h = 3, 32
    dy64 = nhwc_bf16(n, h, h, 64, 125)
    add = nhwc_bf16(n, h, h, 64, 126)
    wd = ops.pack_weight_dgrad(weight(64, 64, 3, 127))
    whole, dx, pad = _guarded((n, h, h, 64), torch.bfloat16, S)
    ops.conv_dgrad(ops.conv_desc(n, h, h, 64, 64, 3, 1, flags=128), dy64, wd, dx, add)
    torch.cuda.synchronize()
    ref = torch.nn.grad.conv2d_input((n, 64, h, h), weight(64, 64, 3, 127), to_nchw(dy64), padding=1) + to_nchw(add)
    assert _guards_intact(whole, pad, S) and rel_err(to_nchw(dx), ref) < BF16_TOL
    # two-map 1x1 conv
    n, h, c, co = 5, 4, 128, 128
    x0, x1 = nhwc_bf16(n, h, h, c, 128), nhwc_bf16(n, h, h, c, 129)
    w2 = (torch.randn(2, co, c, device="cuda") * 0.05).to(torch.bfloat16)
    whole, y, pad = _guarded((n, h, h, co), torch.bfloat16, S)
    ops.conv1x1_cat_fprop(ops.conv_desc(n, h, h, c, co, 1, 1), x0, x1, w2, y)
    torch.cuda.synchronize()
    assert _guards_intact(whole, pad, S) and torch.isfinite(y.float()).all()
    # weight re-layout: both operand arenas
    g = torch.Generator(device="cuda").manual_seed(130)
    shapes = [(96, 24, 1), (64, 64, 3), (40, 36, 3)]
    entries, masters, offs, wf_n, wd_n = [], [], 0, 0, 0
    for co_, ci_, k in shapes:
        w = torch.randn(co_, ci_, k, k, device="cuda", generator=g)
        masters.append(w)
        cip, cop = (ci_ + 63) // 64 * 64, (co_ + 63) // 64 * 64
        entries.append(dict(src_off=offs, wf_off=wf_n, wd_off=wd_n, cout=co_, cin=ci_, taps=k * k, cin_pad=cip, cout_pad=cop,
                            kflat=2))
        offs += w.numel(); wf_n += cop * cip * k * k; wd_n += cop * cip * k * k
    master = torch.cat([w.permute(2, 3, 1, 0).reshape(-1) for w in masters]).contiguous()
    whole_f, wf, pad_f = _guarded((wf_n,), torch.bfloat16, S)
    whole_d, wdd, pad_d = _guarded((wd_n,), torch.bfloat16, S)
    wf.zero_(); wdd.zero_()
    ops.repack_weights(master, wf, wdd, ops.make_repack_table(entries, "cuda"), len(entries))
    torch.cuda.synchronize()
    assert _guards_intact(whole_f, pad_f, S) and _guards_intact(whole_d, pad_d, S)
    for e, w in zip(entries, masters):
        kk, cop, cip = e["taps"], e["cout_pad"], e["cin_pad"]
        f = wf[e["wf_off"]:e["wf_off"] + kk * cop * cip].view(kk, cop, cip)
        d = wdd[e["wd_off"]:e["wd_off"] + kk * cop * cip].view(kk, cip, cop)
        assert torch.equal(f[:, :e["cout"], :e["cin"]], ops.pack_weight_fprop(w))
        assert torch.equal(d[:, :e["cin"], :e["cout"]], ops.pack_weight_dgrad(w))
    # BatchNorm apply with fused finalize, channel count that needs a rounded grid, rows not a multiple of anything
    C, rows = 192, 1001
    yv = (torch.randn(rows, C, device="cuda", generator=g) * 2).to(torch.bfloat16)
    o = {k: torch.empty(C, device="cuda") for k in ("scale", "shift", "mean", "rstd")}
    rec = ops.bn_stats(yv.float().sum(0).contiguous(), (yv.float() ** 2).sum(0).contiguous(),
                       torch.ones(C, device="cuda"), torch.zeros(C, device="cuda"), rows, 1e-5, 0.1, None, None,
                       o["scale"], o["shift"], o["mean"], o["rstd"])
    whole, out, pad = _guarded((rows, C), torch.bfloat16, S)
    ops.bn_act_apply_stats(yv, rec, out, rows, C, relu=True)
    whole2, out2, pad2 = _guarded((rows, C), torch.bfloat16, S)
    ops.bn_act_apply(yv, o["scale"], o["shift"], out2, rows, C, relu=True)
    torch.cuda.synchronize()
    assert _guards_intact(whole, pad, S) and _guards_intact(whole2, pad2, S) and torch.equal(out, out2)
    assert rel_err(out.float(), F.relu(F.batch_norm(yv.float(), None, None, training=True))) < BF16_TOL


@pytest.mark.parametrize("shape", [(5, 32, 32, 64, 1), (3, 16, 16, 192, 1), (7, 8, 8, 128, 1), (4, 32, 32, 64, 2),
                                   (6, 16, 16, 320, 2), (2, 4, 4, 8, 1)])
def test_maxpool3x3(ops, shape):
    """b2c_maxpool3x3_fwd / _bwd = nn.MaxPool2d(3, stride, padding=1) (models/googlenet.py:40,67): identical outputs and
    input gradients on identical bf16 inputs, ties (many after a ReLU) resolved like torch; the optional `add`."""
    n, h, w, c, stride = shape
    g = torch.Generator(device="cuda").manual_seed(160)
    x = torch.randn(n, h, w, c, device="cuda", generator=g).clamp_min(0).to(torch.bfloat16)       # many ties at 0
    ho, wo = h // stride, w // stride
    _, y, _ = _guarded((n, ho, wo, c), torch.bfloat16, 3.0)
    arg = torch.empty(n, ho, wo, c, device="cuda", dtype=torch.uint8)
    ops.maxpool3x3_fwd(x, y, arg, stride)
    xr = to_nchw(x).requires_grad_()
    yr = F.max_pool2d(xr, 3, stride, 1)
    dy = nhwc_bf16(n, ho, wo, c, 161)
    yr.backward(to_nchw(dy))
    add = nhwc_bf16(n, h, w, c, 162)
    whole, dx, pad = _guarded((n, h, w, c), torch.bfloat16, 3.0)
    ops.maxpool3x3_bwd(dy, arg, dx, stride)
    dx2 = torch.empty_like(dx)
    ops.maxpool3x3_bwd(dy, arg, dx2, stride, add=add)
    torch.cuda.synchronize()
    assert torch.equal(to_nchw(y), yr.detach())
    assert _guards_intact(whole, pad, 3.0)
    # sums of up to 9 bf16 values are rounded once here and accumulated in fp32 by torch: compare at bf16 resolution
    assert rel_err(to_nchw(dx), xr.grad) < 3e-3 and (to_nchw(dx) != 0).eq(xr.grad != 0).all()
    assert rel_err(to_nchw(dx2), xr.grad + to_nchw(add)) < 3e-3


def test_copy_channels(ops):
    g = torch.Generator(device="cuda").manual_seed(163)
    src = torch.randn(37, 8, 192, device="cuda", generator=g).to(torch.bfloat16)
    whole, dst, pad = _guarded((37, 8, 320), torch.bfloat16, 9.0)
    dst.fill_(1.0)
    ops.copy_channels(src, 64, dst, 200, 104)
    torch.cuda.synchronize()
    assert _guards_intact(whole, pad, 9.0)
    assert torch.equal(dst[..., 200:304], src[..., 64:168]) and (dst[..., :200] == 1).all() and (dst[..., 304:] == 1).all()


@pytest.mark.parametrize("shape", [(6, 16, 16, (64, 128, 64, 64), 128), (5, 8, 8, (384, 192, 64, 128), 256),
                                   (3, 32, 32, (64, 128), 64), (130, 8, 8, (256, 320, 128, 128), 384)])
def test_conv1x1_over_a_concat_of_different_widths(ops, shape):
    """b2c_conv1x1_catr_fprop: the 1x1 convs of an Inception module over torch.cat([y1, y2, y3, y4], 1) of the previous
    one (models/googlenet.py:52-57), every source through its own tensor map with its own number of K chunks."""
    n, h, w, cs, cout = shape
    xs = [nhwc_bf16(n, h, w, c, 170 + t) for t, c in enumerate(cs)]
    wt = weight(cout, sum(cs), 1, 175)
    ref = F.conv2d(torch.cat([to_nchw(x) for x in xs], 1), wt)
    cmax = max(cs)
    wn = torch.zeros(len(cs), cout, cmax, device="cuda", dtype=torch.bfloat16)
    off = 0
    for t, c in enumerate(cs):
        wn[t, :, :c] = wt[:, off:off + c, 0, 0].to(torch.bfloat16)
        off += c
    whole, y, pad = _guarded((n, h, w, cout), torch.bfloat16, 555.0)
    ps = torch.zeros(ops.MAX_STAT_ROWS, cout, device="cuda")
    pq = torch.zeros(ops.MAX_STAT_ROWS, cout, device="cuda")
    rows = ops.conv1x1_catr_fprop(ops.conv_desc(n, h, w, cmax, cout, 1, 1), xs, wn, y, ps, pq)
    torch.cuda.synchronize()
    assert _guards_intact(whole, pad, 555.0)
    assert torch.isfinite(to_nchw(y)).all() and rel_err(to_nchw(y), ref) < BF16_TOL
    assert rel_err(pq[:rows].sum(0), (ref * ref).sum((0, 2, 3))) < F32_TOL


@pytest.mark.parametrize("shape", [(6, 32, 128, 2, 1), (5, 16, 256, 4, 1), (9, 8, 512, 8, 1), (4, 16, 256, 2, 2),
                                   (3, 32, 128, 2, 2), (7, 8, 256, 2, 1)])
def test_grouped_conv_on_channel_slices(ops, shape):
    """B2cConvDesc.x_ld / y_ld: one group of nn.Conv2d(C, C, 3, stride, padding=1, groups=g) (models/resnext.py:19) is
    the dense C/g -> C/g conv on a channel slice of the C-wide tensors — forward with statistics accumulated into the
    slice of one row, data gradient into a slice of dx (nothing outside the slice is written), weight gradient from
    the two slices — against fp32 torch on the same bf16 inputs."""
    n, hw, C, g, stride = shape
    cg, ho = C // g, hw // stride
    x, dy = nhwc_bf16(n, hw, hw, C, 180), nhwc_bf16(n, ho, ho, C, 181)
    w = weight(C, cg, 3, 182)
    y = torch.full((n, ho, ho, C), 7.0, device="cuda", dtype=torch.bfloat16)
    dx = torch.full((n, hw, hw, C), 7.0, device="cuda", dtype=torch.bfloat16)
    stats = torch.zeros(2, C, device="cuda")
    dw = torch.zeros(C, cg, 3, 3, device="cuda")
    scratch = torch.zeros(ops.WGRAD_SCRATCH_FLOATS, device="cuda")
    d = ops.conv_desc(n, hw, hw, cg, cg, 3, stride, flags=128 | 512, x_ld=C, y_ld=C)
    dwd = ops.conv_desc(n, hw, hw, cg, cg, 3, stride, x_ld=C, y_ld=C)
    for j in (g - 1, 0):                           # all groups but one: the untouched slice must stay untouched
        wj = w[j * cg:(j + 1) * cg]
        xs, ys = x.view(-1)[j * cg:], y.view(-1)[j * cg:]
        ops.conv_fprop(d, xs, ops.pack_weight_fprop(wj), ys, stats[0, j * cg:], stats[1, j * cg:])
        ops.conv_dgrad(d, dy.view(-1)[j * cg:], ops.pack_weight_dgrad(wj), dx.view(-1)[j * cg:])
        ops.conv_wgrad(dwd, xs, dy.view(-1)[j * cg:], dw[j * cg:(j + 1) * cg], scratch=scratch)
    torch.cuda.synchronize()
    xr, wr = to_nchw(x).requires_grad_(), w.clone().requires_grad_()
    yr = F.conv2d(xr, wr, stride=stride, padding=1, groups=g)
    yr.backward(to_nchw(dy))
    for j in (g - 1, 0):
        sl = slice(j * cg, (j + 1) * cg)
        assert rel_err(to_nchw(y)[:, sl], yr.detach()[:, sl]) < BF16_TOL
        assert rel_err(to_nchw(dx)[:, sl], xr.grad[:, sl]) < BF16_TOL
        assert rel_err(dw[sl], wr.grad[sl]) < 2e-3
        assert rel_err(stats[1, sl], (yr.detach()[:, sl] ** 2).sum((0, 2, 3))) < F32_TOL
    if g > 2:
        mid = slice(cg, (g - 1) * cg)
        assert (y[..., mid] == 7.0).all() and (dx[..., mid] == 7.0).all() and (stats[:, mid] == 0).all()
        assert (dw[mid] == 0).all()
