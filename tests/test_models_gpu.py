"""GPU parity for the other model families of BASELINE.json (config 3: VGG-16, config 4: PreActResNet-18) and
the kernels only they use (max-pool, residual add in the conv epilogue, identity-bypass add in BN backward).

New kernels: tight against fp32 torch on identical inputs. Whole models: forward against the fp32 oracle
(oracle/nets.py, pinned to the reference in tests/test_oracle.py); gradients end to end are compared the way
tests/test_network_gpu.py explains — the layers nearest the loss tightly, every tensor by direction (cosine),
and the overall error level against stock torch bf16 autocast on the same inputs.
"""
import json
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def rel(a, b):
    a, b = a.double(), b.double()
    return ((a - b).norm() / b.norm().clamp_min(1e-30)).item()


def cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return (a @ b / (a.norm() * b.norm()).clamp_min(1e-30)).item()


@pytest.fixture(scope="module")
def ops():
    import b200cifar
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    return b200cifar.ops


def test_maxpool_fwd_bwd(ops):
    g = torch.Generator(device="cuda").manual_seed(50)
    for (n, h, w, c) in [(4, 32, 32, 64), (3, 4, 4, 512), (2, 2, 2, 512)]:
        x = torch.randn(n, h, w, c, device="cuda", generator=g).clamp_min(0).to(torch.bfloat16)   # many ties at 0
        y = torch.empty(n, h // 2, w // 2, c, device="cuda", dtype=torch.bfloat16)
        ops.maxpool2x2_fwd(x, y)
        xr = x.float().permute(0, 3, 1, 2).requires_grad_()
        yr = F.max_pool2d(xr, 2, 2)
        dy = torch.randn(n, h // 2, w // 2, c, device="cuda", generator=g).to(torch.bfloat16)
        yr.backward(dy.float().permute(0, 3, 1, 2))
        dx = torch.empty_like(x)
        ops.maxpool2x2_bwd(x, y, dy, dx)
        torch.cuda.synchronize()
        assert torch.equal(y.float().permute(0, 3, 1, 2), yr.detach())
        assert torch.equal(dx.float().permute(0, 3, 1, 2), xr.grad)


def test_conv_fprop_residual_add_and_stats(ops):
    n, h, w, cin, cout = 8, 16, 16, 128, 128
    g = torch.Generator(device="cuda").manual_seed(51)
    x = torch.randn(n, h, w, cin, device="cuda", generator=g).to(torch.bfloat16)
    add = torch.randn(n, h, w, cout, device="cuda", generator=g).to(torch.bfloat16)
    wt = (torch.randn(cout, cin, 3, 3, device="cuda", generator=g) * 0.03).to(torch.bfloat16).float()
    for flags in (0, 1):                       # halo kernel and per-tap kernel
        y = torch.empty(n, h, w, cout, device="cuda", dtype=torch.bfloat16)
        ps = torch.zeros(ops.MAX_STAT_ROWS, cout, device="cuda")
        pq = torch.zeros(ops.MAX_STAT_ROWS, cout, device="cuda")
        rows = ops.conv_fprop(ops.conv_desc(n, h, w, cin, cout, 3, 1, flags=flags), x, ops.pack_weight_fprop(wt), y,
                              ps, pq, add_src=add)
        torch.cuda.synchronize()
        ref = F.conv2d(x.float().permute(0, 3, 1, 2), wt, padding=1) + add.float().permute(0, 3, 1, 2)
        assert rel(y.float().permute(0, 3, 1, 2), ref) < 1e-2
        assert rel(pq[:rows].sum(0), (ref * ref).sum((0, 2, 3))) < 1e-3     # statistics include the add


def test_bn_bwd_apply_with_bypass_add(ops):
    rows, C = 2048, 128
    g = torch.Generator(device="cuda").manual_seed(52)
    dz = torch.randn(rows, C, device="cuda", generator=g).to(torch.bfloat16)
    y = torch.randn(rows, C, device="cuda", generator=g).to(torch.bfloat16)
    add = torch.randn(rows, C, device="cuda", generator=g).to(torch.bfloat16)
    ca, cb, cc = (torch.randn(C, device="cuda", generator=g) for _ in range(3))
    dy = torch.empty_like(dz)
    ops.bn_bwd_apply(dz, y, ca, cb, cc, dy, rows, C, add=add)
    torch.cuda.synchronize()
    assert rel(dy.float(), ca * dz.float() + cb * y.float() + cc + add.float()) < 1e-2


@pytest.mark.parametrize("shape", [(4, 32, 32, 96, 96, 1), (4, 32, 32, 192, 144, 2), (8, 8, 8, 384, 384, 1),
                                   (8, 8, 8, 576, 576, 2), (16, 4, 4, 960, 960, 1), (2, 32, 32, 64, 32, 1),
                                   (5, 16, 16, 192, 144, 1), (3, 16, 16, 192, 192, 2), (40, 8, 8, 64, 64, 1)])
def test_depthwise_conv(ops, shape):
    n, h, w, cpad, c, stride = shape
    g = torch.Generator(device="cuda").manual_seed(60)
    x = torch.zeros(n, h, w, cpad, device="cuda", dtype=torch.bfloat16)
    x[..., :c] = torch.randn(n, h, w, c, device="cuda", generator=g).to(torch.bfloat16)
    wt = torch.randn(c, 1, 3, 3, device="cuda", generator=g) * 0.3
    ho, wo = h // stride, w // stride
    y = torch.full((n, ho, wo, cpad), float("nan"), device="cuda", dtype=torch.bfloat16)
    ps = torch.zeros(ops.MAX_DW_STAT_ROWS, cpad, device="cuda")
    pq = torch.zeros(ops.MAX_DW_STAT_ROWS, cpad, device="cuda")
    rows = ops.dwconv_fprop(x, wt, y, c, stride, ps, pq)
    xr = x[..., :c].float().permute(0, 3, 1, 2).requires_grad_()
    wr = wt.clone().requires_grad_()
    yr = F.conv2d(xr, wr, stride=stride, padding=1, groups=c)
    dy = torch.zeros(n, ho, wo, cpad, device="cuda", dtype=torch.bfloat16)
    dy[..., :c] = torch.randn(n, ho, wo, c, device="cuda", generator=g).to(torch.bfloat16)
    yr.backward(dy[..., :c].float().permute(0, 3, 1, 2))
    dx = torch.full((n, h, w, cpad), float("nan"), device="cuda", dtype=torch.bfloat16)
    ops.dwconv_dgrad(dy, wt, dx, c, stride)
    dw = torch.zeros(c, 1, 3, 3, device="cuda")
    ops.dwconv_wgrad(x, dy, dw, c, stride)
    torch.cuda.synchronize()
    assert rel(y[..., :c].float().permute(0, 3, 1, 2), yr.detach()) < 1e-2
    assert (y[..., c:] == 0).all() and (dx[..., c:] == 0).all()            # padded channels stay zero
    assert rel(ps[:rows].sum(0)[:c], yr.detach().sum((0, 2, 3))) < 5e-3
    assert rel(pq[:rows].sum(0)[:c], (yr.detach() ** 2).sum((0, 2, 3))) < 1e-3
    assert rel(dx[..., :c].float().permute(0, 3, 1, 2), xr.grad) < 1e-2
    assert rel(dw, wr.grad) < 1e-3
    # statistics accumulated into one row (b2c_dwconv_fprop_acc): same output, same sums
    y2 = torch.full_like(y, float("nan"))
    acc = torch.zeros(2, cpad, device="cuda")
    ops.dwconv_fprop_acc(x, wt, y2, c, stride, acc[0], acc[1])
    torch.cuda.synchronize()
    assert torch.equal(y, y2)
    assert (acc[0] - ps[:rows].sum(0)).abs().max() <= 1e-4 * ps[:rows].abs().sum(0).max() + 1e-3
    assert rel(acc[1], pq[:rows].sum(0)) < 1e-5


def test_pointwise_conv_with_padded_channels(ops):
    """1x1 conv 24 -> 144 in the 64-padded layout (64 -> 192): padded weights are zero, wgrad masks them."""
    n, hw, cin, cout, cin_p, cout_p = 4, 32, 24, 144, 64, 192
    g = torch.Generator(device="cuda").manual_seed(61)
    x = torch.zeros(n, hw, hw, cin_p, device="cuda", dtype=torch.bfloat16)
    x[..., :cin] = torch.randn(n, hw, hw, cin, device="cuda", generator=g).to(torch.bfloat16)
    wt = (torch.randn(cout, cin, 1, 1, device="cuda", generator=g) * 0.2)
    master = wt.flatten().contiguous()
    wf = torch.zeros(cout_p * cin_p, device="cuda", dtype=torch.bfloat16)
    wd = torch.zeros(cout_p * cin_p, device="cuda", dtype=torch.bfloat16)
    table = ops.make_repack_table([dict(src_off=0, wf_off=0, wd_off=0, cout=cout, cin=cin, taps=1, cin_pad=cin_p,
                                        cout_pad=cout_p)], "cuda")
    ops.repack_weights(master, wf, wd, table, 1)
    y = torch.empty(n, hw, hw, cout_p, device="cuda", dtype=torch.bfloat16)
    d = ops.conv_desc(n, hw, hw, cin_p, cout_p, 1, 1, cin_valid=cin, cout_valid=cout)
    ops.conv_fprop(d, x, wf.view(1, cout_p, cin_p), y)
    wq = wt.to(torch.bfloat16).float()
    ref = F.conv2d(x[..., :cin].float().permute(0, 3, 1, 2), wq)
    dy = torch.zeros(n, hw, hw, cout_p, device="cuda", dtype=torch.bfloat16)
    dy[..., :cout] = torch.randn(n, hw, hw, cout, device="cuda", generator=g).to(torch.bfloat16)
    dx = torch.empty(n, hw, hw, cin_p, device="cuda", dtype=torch.bfloat16)
    ops.conv_dgrad(d, dy, wd.view(1, cin_p, cout_p), dx)
    dw = torch.zeros(cout, cin, 1, 1, device="cuda")
    ops.conv_wgrad(d, x, dy, dw)
    torch.cuda.synchronize()
    assert rel(y[..., :cout].float().permute(0, 3, 1, 2), ref) < 1e-2 and (y[..., cout:] == 0).all()
    dyr = dy[..., :cout].float().permute(0, 3, 1, 2)
    assert rel(dx[..., :cin].float().permute(0, 3, 1, 2), torch.nn.grad.conv2d_input((n, cin, hw, hw), wq, dyr)) < 1e-2
    assert (dx[..., cin:] == 0).all()
    assert rel(dw, torch.nn.grad.conv2d_weight(x[..., :cin].float().permute(0, 3, 1, 2), (cout, cin, 1, 1), dyr)) < 1e-3


# per-model bounds for the END-TO-END comparison with fp32 (see the module docstring): MobileNetV2 is 52 convs deep
# with 4x4 feature maps at batch 128, where stock torch bf16 autocast itself is 85 % off on the median gradient.
# ResNet-50 (49 convs, 4x expansion) at random init is worse still: stock autocast shows 24 % on the logits and
# > 100 % on the early-layer gradients, and with lr 0.02 the fp32 oracle itself diverges on this batch (the CUDA
# path follows the same diverging trajectory step for step) — so it trains with lr 0.002 here.
BOUNDS = {"vgg16": dict(logits=5e-2, near=6e-2, near_cos=0.99, min_cos=0.7, drop=0.7, lr=0.02),
          "preact18": dict(logits=5e-2, near=6e-2, near_cos=0.99, min_cos=0.7, drop=0.7, lr=0.02),
          "mobilenetv2": dict(logits=0.25, near=0.15, near_cos=0.99, min_cos=None, drop=0.92, lr=0.02),
          "resnet50": dict(logits=0.3, near=0.3, near_cos=0.95, min_cos=None, drop=0.995, lr=0.002),
          # SimpleDLA: 33 convs deep; measured 8.2 % on the logits (profiles/r2_parity/parity_simpledla.json)
          # MobileNet v1: 27 layers without a residual path; at random init the per-layer bf16 noise grows ~1.35x per block
          # (0.3 % after the stem -> 28 % on the logits, profiles/r2_parity/parity_mobilenet.json); stock bf16 autocast
          # shows the same median gradient error (1.04 vs our 1.02)
          "mobilenet": dict(logits=0.45, near=0.3, near_cos=0.95, min_cos=None, drop=0.95, lr=0.02),
          "simpledla": dict(logits=0.12, near=0.12, near_cos=0.99, min_cos=None, drop=0.9, lr=0.02),
          "preact50": dict(logits=0.3, near=0.3, near_cos=0.95, min_cos=None, drop=0.995, lr=0.002),
          "googlenet": dict(logits=0.12, near=0.15, near_cos=0.98, min_cos=None, drop=0.9, lr=0.02),
          # SENet18: measured 0.65 % on the logits (profiles/r2_parity/parity_senet18.json)
          "senet18": dict(logits=0.03, near=0.06, near_cos=0.99, min_cos=0.7, drop=0.9, lr=0.02),
          # DLA: 49 convs deep; measured 9.7 % on the logits, 12.7 % on the last BatchNorm's bias gradient — stock bf16
          # autocast shows 12.9 % on the same tensor (profiles/r2_parity/parity_dla.json)
          "dla": dict(logits=0.15, near=0.2, near_cos=0.98, min_cos=None, drop=0.9, lr=0.02),
          # ResNeXt-29 (29 convs, grouped 3x3, 8x8 final maps): measured 1.6 % / 1.9 % on the logits
          # (profiles/r2_parity/parity_resnext29_*.json); this batch's loss rises before it falls, for the oracle too
          "resnext29_2x64d": dict(logits=0.06, near=0.3, near_cos=0.95, min_cos=None, drop=0.9, lr=0.02),
          # (4 groups run group by group; at lr 0.02 the fp32 oracle's loss on this batch climbs 2.5 -> 22.7 and ours
          # follows it to three digits, profiles/r2_parity/train_curve_resnext29_4x64d_lr0.02.json: lr 0.002 here)
          "resnext29_4x64d": dict(logits=0.06, near=0.3, near_cos=0.95, min_cos=None, drop=0.995, lr=0.002),
          "resnext29_32x4d": dict(logits=0.06, near=0.3, near_cos=0.95, min_cos=None, drop=0.9, lr=0.02),
          # RegNetX (41 / 68 convs; widths 24..384 padded to the 64-channel layout; 3..46 groups per conv)
          # measured 5.1 % / 4.8 % on the logits (profiles/r2_parity/parity_regnetx_*.json). With lr 0.02 the fp32 oracle's
          # loss on this batch climbs from 3.3 to 12 for the 400MF network (ours follows it step for step,
          # profiles/r2_parity/train_curve_regnetx_400mf_lr0.02.json), so both train with lr 0.002 here, like ResNet-50
          "regnetx_200mf": dict(logits=0.12, near=0.3, near_cos=0.95, min_cos=None, drop=0.995, lr=0.002),
          "regnetx_400mf": dict(logits=0.12, near=0.3, near_cos=0.95, min_cos=None, drop=0.995, lr=0.002),
          "regnety_400mf": dict(logits=0.12, near=0.3, near_cos=0.95, min_cos=None, drop=0.995, lr=0.002)}

MODELS = {
    "vgg16": ("vgg_spec", "build_vgg_params", "vgg_forward", "features.0", ["classifier.weight", "features.41.weight",
                                                                              "features.41.bias"]),
    "preact18": ("preact_resnet_spec", "build_preact_params", "preact_forward", "conv1",
                 ["linear.weight", "layer4.1.conv2.weight", "layer4.1.bn2.weight"]),
    "mobilenetv2": ("mobilenetv2_spec", "build_mobilenetv2_params", "mobilenetv2_forward", "conv1",
                    ["linear.weight", "bn2.weight", "bn2.bias"]),
    "resnet50": ("resnet50_spec", "build_bottleneck_params", "bottleneck_forward", "conv1",
                 ["linear.weight", "layer4.2.bn3.weight", "layer4.2.bn3.bias"]),
    "mobilenet": ("mobilenet_spec", "build_mobilenet_params", "mobilenet_forward", "conv1",
                  ["linear.weight", "layers.12.bn2.weight", "layers.12.bn2.bias"]),
    # ResNeXt-29 (models/resnext.py): 1x1 stem, grouped 3x3 convs (2 groups of 64.. / 32 groups of 4..), avg_pool2d(8)
    "resnext29_2x64d": ("resnext_spec", "build_resnext_params", "resnext_forward", "conv1",
                        ["linear.weight", "layer3.2.bn3.weight", "layer3.2.bn3.bias"]),
    "resnext29_4x64d": ("resnext29_4x64d_spec", "build_resnext_4x64d_params", "resnext_4x64d_forward", "conv1",
                        ["linear.weight", "layer3.2.bn3.weight", "layer3.2.bn3.bias"]),
    "resnext29_32x4d": ("resnext29_32x4d_spec", "build_resnext_32x4d_params", "resnext_32x4d_forward", "conv1",
                        ["linear.weight", "layer3.2.bn3.weight", "layer3.2.bn3.bias"]),
    "regnetx_200mf": ("regnetx_spec", "build_regnetx_params", "regnetx_forward", "conv1",
                      ["linear.weight", "layer4.6.bn3.weight", "layer4.6.bn3.bias"]),
    "regnetx_400mf": ("regnetx_400mf_spec", "build_regnetx_400mf_params", "regnetx_400mf_forward", "conv1",
                      ["linear.weight", "layer4.11.bn3.weight", "layer4.11.bn3.bias"]),
    # DLA (models/dla.py): level-2 trees whose Root reads four tensors; two blocks read a tree's input
    "dla": ("dla_spec", "build_dla_params", "dla_forward", "base.0",
            ["linear.weight", "layer6.root.bn.weight", "layer6.root.bn.bias"]),
    "regnety_400mf": ("regnety_400mf_spec", "build_regnety_400mf_params", "regnety_400mf_forward", "conv1",
                      ["linear.weight", "layer4.11.bn3.weight", "layer4.11.bn3.bias"]),
    # GoogLeNet (models/googlenet.py): Inception modules, concatenations read through per-tensor maps, 3x3 max-pools
    "googlenet": ("googlenet_spec", "build_googlenet_params", "googlenet_forward", "pre_layers.0",
                  ["linear.weight", "b5.b1.1.weight", "b5.b4.2.bias"]),
    "preact50": ("preact_resnet50_spec", "build_preact_bottleneck_params", "preact_bottleneck_forward", "conv1",
                 ["linear.weight", "layer4.2.conv3.weight", "layer4.2.bn3.weight"]),
    # SENet18 (models/senet.py): pre-activation blocks with squeeze-and-excitation
    "senet18": ("senet_spec", "build_senet_params", "senet_forward", "conv1",
                ["linear.weight", "layer4.1.fc2.weight", "layer4.1.conv2.weight"]),
    # SimpleDLA, what main.py:71 constructs (models/dla_simple.py): aggregation trees, Root = 1x1 conv over a concat
    "simpledla": ("simple_dla_spec", "build_simpledla_params", "simpledla_forward", "base.0",
                  ["linear.weight", "layer6.root.bn.weight", "layer6.root.bn.bias"]),
}


@pytest.mark.parametrize("model", list(MODELS))
def test_model_training_step(ops, model):
    from b200cifar import engine
    from oracle import nets
    spec_fn, build_fn, fwd_fn, first_conv, near_loss = MODELS[model]
    N = 128
    params, buffers = getattr(nets, build_fn)(0)
    forward = getattr(nets, fwd_fn)
    rs = np.random.RandomState(4)
    x = torch.from_numpy(rs.standard_normal((N, 3, 32, 32)).astype(np.float32)).cuda()
    target = torch.from_numpy(rs.randint(0, 10, size=(N,)).astype(np.int64)).cuda()
    cp = {k: v.cuda() for k, v in params.items()}
    cb = {k: v.cuda() for k, v in buffers.items()}
    taps = {}
    loss_ref, logits_ref, grads_ref = nets.train_step_grads(x, target, cp, cb, forward=forward, taps=taps)
    eng = engine.Engine(getattr(engine, spec_fn)(), N)
    eng.load_params(params, {k: v for k, v in buffers.items() if not k.endswith("num_batches_tracked")})
    ops.stem_im2col(x, eng.x0, taps=eng.stem_taps)
    eng.zero_grad()
    eng.forward(True)
    loss = torch.zeros(1, device="cuda")
    ops.ce_loss(eng.logits, target, N, 10, eng.dlogits, loss)
    eng.backward()
    torch.cuda.synchronize()
    def ref_conv_out(name):
        t = taps[name].detach()
        b = cp.get(name + ".bias")      # VGG convs carry a bias that the following BatchNorm cancels; the CUDA path
        return t if b is None else t - b.view(1, -1, 1, 1)     # leaves it out of the stored conv output
    fwd_err = {u.name: rel(u.y[..., :taps[u.name].shape[1]].float().permute(0, 3, 1, 2), ref_conv_out(u.name))
               for u in eng.units if u.name in taps}
    grads = eng.named_grads()
    # a conv bias followed by BatchNorm has a mathematically zero gradient (fp32 shows round-off noise, we write 0)
    # (not the squeeze-and-excitation 1x1 convs: their bias feeds a ReLU / sigmoid)
    skip = {k for k in grads_ref if k.endswith(".bias") and cp[k[:-5] + ".weight"].dim() == 4
            and (k[:-5] + ".weight") not in eng.spec.get("se_fc", ())}
    assert all(grads[k].abs().max().item() == 0 for k in skip)
    keep = {k: v for k, v in grads_ref.items() if k not in skip and v.norm().item() > 1e-6}
    g_rel = {k: rel(grads[k], v) for k, v in keep.items()}
    g_cos = {k: cos(grads[k], v) for k, v in keep.items()}
    # stock bf16 autocast for context
    leaves = {k: v.detach().clone().requires_grad_(True) for k, v in cp.items()}
    ab = {k: v.clone().cuda() for k, v in buffers.items()}
    with torch.autocast("cuda", dtype=torch.bfloat16):
        ls = F.cross_entropy(forward(x, leaves, ab, training=True).float(), target)
    ag = torch.autograd.grad(ls, list(leaves.values()))
    a_rel = {k: rel(g.float(), grads_ref[k]) for k, g in zip(leaves, ag) if k in keep}
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(dict(forward=fwd_err, logits=rel(eng.logits, logits_ref), grad_rel=g_rel, grad_cos=g_cos,
                   autocast_grad_rel=a_rel), open(f"gpurun_out/parity_{model}.json", "w"), indent=1)
    assert fwd_err[first_conv] < 1e-2                                   # first layer: no compounding yet
    bnd = BOUNDS[model]
    assert rel(eng.logits, logits_ref) < bnd["logits"] and abs(loss.item() - loss_ref.item()) < 2e-2
    for k in near_loss:                                                 # nearest the loss: tight
        assert g_rel[k] < bnd["near"] and g_cos[k] > bnd["near_cos"], (k, g_rel[k], g_cos[k])
    if bnd["min_cos"] is not None:
        assert min(g_cos.values()) > bnd["min_cos"], sorted(g_cos.items(), key=lambda kv: kv[1])[:5]
    assert np.median(list(g_rel.values())) < 1.5 * np.median(list(a_rel.values())) + 1e-2
    for name, b in eng.named_buffers().items():                         # BN running statistics
        if name.endswith("running_var"):
            assert rel(b, cb[name]) < 3e-2, name


@pytest.mark.parametrize("model", list(MODELS))
def test_model_trains(ops, model):
    """A few SGD steps on one batch through the Trainer (CUDA graph): the loss must fall like the fp32 oracle's."""
    from b200cifar import data, engine, trainer
    from oracle import nets
    spec_fn, build_fn, fwd_fn, _, _ = MODELS[model]
    N = 128
    params, buffers = getattr(nets, build_fn)(0)
    forward = getattr(nets, fwd_fn)
    images, labels = data.learnable_cifar(512, seed=3)
    lr = BOUNDS[model]["lr"]
    tr = trainer.Trainer(getattr(engine, spec_fn)(), N, lr=lr)
    tr.load_params(params, {k: v for k, v in buffers.items() if not k.endswith("num_batches_tracked")})
    tr.set_dataset(images, labels)
    idx = np.arange(N)
    ijf = np.tile(np.array([[4, 4, 0]], dtype=np.int8), (N, 1))        # identity crop, no flip
    losses = []
    for _ in range(8):
        tr.step(idx, ijf)
        losses.append(tr.read_loss())
    # fp32 oracle on the same batch
    from oracle import augment as oaug
    xb = torch.from_numpy(oaug.normalize_only(images[:N])).cuda()
    tb = torch.from_numpy(labels[:N]).cuda()
    leaves = {k: v.detach().clone().cuda().requires_grad_(True) for k, v in params.items()}
    cb = {k: v.clone().cuda() for k, v in buffers.items()}
    opt = torch.optim.SGD(list(leaves.values()), lr=lr, momentum=0.9, weight_decay=5e-4)
    ref = []
    for _ in range(8):
        opt.zero_grad()
        l = F.cross_entropy(forward(xb, leaves, cb, training=True), tb)
        l.backward()
        opt.step()
        ref.append(l.item())
    json.dump(dict(ours=losses, oracle=ref), open(f"gpurun_out/train_curve_{model}.json", "w"))
    assert abs(losses[0] - ref[0]) < 3e-2
    assert losses[-1] < BOUNDS[model]["drop"] * losses[0] and abs(losses[-1] - ref[-1]) < 0.35 * ref[0]
    assert max(abs(a - b) for a, b in zip(losses[:4], ref[:4])) < 0.05 * ref[0]      # same early trajectory


def test_prefetched_host_batches_match_resident_dataset(ops):
    """The three input paths of the Trainer — dataset resident in HBM (step), host batch copied inside the call
    (step_host_batch) and host batch staged one step ahead on a copy stream (prefetch_host_batch +
    step_prefetched) — see the same samples, crops and labels: identical first-step loss, same trajectory."""
    from b200cifar import data, engine, trainer
    from oracle import nets
    N, steps = 128, 4
    params, buffers = nets.build_resnet_params(0)
    bufs = {k: v for k, v in buffers.items() if not k.endswith("num_batches_tracked")}
    images, labels = data.learnable_cifar(1024, seed=5)
    rng = data.TorchCpuMT19937(11)
    perm = data.randperm(rng, 1024)
    idx = [perm[k * N:(k + 1) * N] for k in range(steps)]
    ijf = [data.draw_crop_flip(rng, N).reshape(N, 3) for _ in range(steps)]
    out = {}
    for mode in ("resident", "host", "prefetch"):
        tr = trainer.Trainer(engine.resnet_spec(), N, lr=0.02, resident_dataset=(mode == "resident"))
        tr.load_params(params, bufs)
        losses = []
        if mode == "resident":
            tr.set_dataset(images, labels)
            for k in range(steps):
                tr.step(idx[k], ijf[k]); losses.append(tr.read_loss())
        elif mode == "host":
            for k in range(steps):
                tr.step_host_batch(images[idx[k]], labels[idx[k]], ijf[k]); losses.append(tr.read_loss())
        else:
            tr.prefetch_host_batch(images[idx[0]], labels[idx[0]], ijf[0])
            for k in range(steps):
                tr.step_prefetched()
                if k + 1 < steps:
                    tr.prefetch_host_batch(images[idx[k + 1]], labels[idx[k + 1]], ijf[k + 1])
                losses.append(tr.read_loss())
        out[mode] = losses
        del tr
    assert out["resident"][0] == out["host"][0] == out["prefetch"][0]
    for k in range(steps):     # later steps differ only by the summation order of the fp32 wgrad atomics
        assert abs(out["host"][k] - out["resident"][k]) < 2e-2 * out["resident"][0]
        assert abs(out["prefetch"][k] - out["resident"][k]) < 2e-2 * out["resident"][0]


@pytest.mark.parametrize("shape", [(4, 32, 32, 128, 96), (8, 16, 16, 192, 192), (16, 8, 8, 384, 384), (32, 4, 4, 960, 960),
                                   (3, 8, 8, 64, 64)])
def test_depthwise_conv_fed_by_the_previous_batchnorm(ops, shape):
    """b2c_dwconv_fprop_bnrelu / b2c_dwconv_wgrad_bnrelu: the depthwise conv reads the previous conv's raw output and
    applies relu(y * scale + shift) while loading (models/mobilenetv2.py:32-33), zero padding after the transform.
    Against fp32 torch on the materialised activation, and against the two-pass path (b2c_bn_act_apply, then
    b2c_dwconv_fprop / wgrad), which differs only by the bf16 rounding of the intermediate it stores."""
    n, h, w, cpad, c = shape
    g = torch.Generator(device="cuda").manual_seed(70)
    yprev = torch.zeros(n, h, w, cpad, device="cuda", dtype=torch.bfloat16)
    yprev[..., :c] = (torch.randn(n, h, w, c, device="cuda", generator=g) * 1.5).to(torch.bfloat16)
    scale = torch.zeros(cpad, device="cuda"); shift = torch.zeros(cpad, device="cuda")
    scale[:c] = torch.rand(c, device="cuda", generator=g) + 0.5
    shift[:c] = torch.randn(c, device="cuda", generator=g) * 0.5          # relu(shift) != 0: padding must not see it
    wt = torch.randn(c, 1, 3, 3, device="cuda", generator=g) * 0.3
    assert ops.dwconv_input_transform_ok(h, w, cpad, 1) and not ops.dwconv_input_transform_ok(h, w, cpad, 2)
    a_ref = F.relu(yprev[..., :c].float() * scale[:c] + shift[:c]).permute(0, 3, 1, 2)
    y_ref = F.conv2d(a_ref, wt, padding=1, groups=c)
    dy = torch.zeros(n, h, w, cpad, device="cuda", dtype=torch.bfloat16)
    dy[..., :c] = torch.randn(n, h, w, c, device="cuda", generator=g).to(torch.bfloat16)
    dw_ref = torch.nn.grad.conv2d_weight(a_ref, (c, 1, 3, 3), dy[..., :c].float().permute(0, 3, 1, 2), padding=1, groups=c)
    # fused
    y = torch.full((n, h, w, cpad), float("nan"), device="cuda", dtype=torch.bfloat16)
    ps = torch.zeros(ops.MAX_DW_STAT_ROWS, cpad, device="cuda"); pq = torch.zeros_like(ps)
    rows = ops.dwconv_fprop_bnrelu(yprev, scale, shift, wt, y, c, ps, pq)
    dw = torch.zeros(c, 1, 3, 3, device="cuda")
    ops.dwconv_wgrad_bnrelu(yprev, scale, shift, dy, dw, c)
    # two passes
    a = torch.empty_like(yprev)
    ops.bn_act_apply(yprev, scale, shift, a, n * h * w, cpad, relu=True)
    y2 = torch.empty_like(y)
    ops.dwconv_fprop(a, wt, y2, c, 1)
    dw2 = torch.zeros(c, 1, 3, 3, device="cuda")
    ops.dwconv_wgrad(a, dy, dw2, c, 1)
    torch.cuda.synchronize()
    assert torch.isfinite(y.float()).all() and (y[..., c:] == 0).all()
    assert rel(y[..., :c].float().permute(0, 3, 1, 2), y_ref) < 1e-2
    assert rel(y.float(), y2.float()) < 1e-2
    assert rel(ps[:rows].sum(0)[:c], y_ref.sum((0, 2, 3))) < 5e-3 or (ps[:rows].sum(0)[:c] - y_ref.sum((0, 2, 3))).abs().max() < 0.5
    assert rel(pq[:rows].sum(0)[:c], (y_ref ** 2).sum((0, 2, 3))) < 2e-3
    assert rel(dw, dw_ref) < 2e-3 and rel(dw, dw2) < 1e-2
    acc = torch.zeros(2, cpad, device="cuda")
    assert ops.dwconv_fprop_bnrelu(yprev, scale, shift, wt, y2, c, acc[0], acc[1], accumulate=True) == 1
    torch.cuda.synchronize()
    assert torch.equal(y, y2) and rel(acc[1], pq[:rows].sum(0)) < 1e-5


@pytest.mark.parametrize("shape", [(6, 32, 128, 2, 1), (5, 16, 256, 4, 2), (9, 8, 512, 32, 1), (4, 16, 128, 32, 1),
                                   (3, 32, 256, 8, 2)])
def test_grouped_conv_through_block_diagonal_operands(ops, shape):
    """nn.Conv2d(C, C, 3, stride, padding=1, groups=g) (models/resnext.py:19) as the C ABI runs it: b2c_repack_weights
    with one table entry per group fills the diagonal blocks of the dense operands, b2c_conv_fprop / dgrad / wgrad run
    the dense conv, b2c_grouped_wgrad_extract folds the dense weight gradient back into the [C][C/g][3][3] parameter
    gradient (accumulating). Against fp32 torch on the same bf16-rounded inputs."""
    n, hw, C, g, stride = shape
    gen = torch.Generator(device="cuda").manual_seed(80 + g)
    cg = C // g
    w = torch.randn(C, cg, 3, 3, device="cuda", generator=gen) * (2.0 / (9 * cg)) ** 0.5
    entries = [dict(src_off=j * cg * cg * 9, wf_off=j * cg * C + j * cg, wd_off=j * cg * C + j * cg, cout=cg, cin=cg,
                    taps=9, cin_pad=C, cout_pad=C, kflat=0) for j in range(g)]
    wf = torch.zeros(9, C, C, device="cuda", dtype=torch.bfloat16)
    wd = torch.zeros(9, C, C, device="cuda", dtype=torch.bfloat16)
    ops.repack_weights(w.reshape(-1).contiguous(), wf, wd, ops.make_repack_table(entries, "cuda"), g)
    wq = w.to(torch.bfloat16).float()
    x = torch.randn(n, hw, hw, C, device="cuda", generator=gen).to(torch.bfloat16)
    ho = hw // stride
    dy = torch.randn(n, ho, ho, C, device="cuda", generator=gen).to(torch.bfloat16)
    y = torch.empty(n, ho, ho, C, device="cuda", dtype=torch.bfloat16)
    dx = torch.empty_like(x)
    d = ops.conv_desc(n, hw, hw, C, C, 3, stride, flags=128)
    ops.conv_fprop(d, x, wf, y)
    ops.conv_dgrad(d, dy, wd, dx)
    dense = torch.zeros(9 * C * C, device="cuda")
    scratch = torch.zeros(ops.WGRAD_SCRATCH_FLOATS, device="cuda")
    ops.conv_wgrad(ops.conv_desc(n, hw, hw, C, C, 3, stride, flags=32 | 128), x, dy, dense, scratch=scratch)
    dw = torch.full((C, cg, 3, 3), 0.5, device="cuda")                    # the extract accumulates
    ops.grouped_wgrad_extract(dense, dw.view(-1), C, C, 9, g)
    torch.cuda.synchronize()
    xr = x.float().permute(0, 3, 1, 2).requires_grad_()
    wr = wq.clone().requires_grad_()
    yr = F.conv2d(xr, wr, stride=stride, padding=1, groups=g)
    yr.backward(dy.float().permute(0, 3, 1, 2))
    # operands: exactly the bf16 weights on the diagonal blocks, zero elsewhere
    dense_w = torch.zeros(C, C, 3, 3, device="cuda")
    for j in range(g):
        dense_w[j * cg:(j + 1) * cg, j * cg:(j + 1) * cg] = wq[j * cg:(j + 1) * cg]
    assert torch.equal(wf.float(), ops.pack_weight_fprop(dense_w).float())
    assert torch.equal(wd.float(), ops.pack_weight_dgrad(dense_w).float())
    assert rel(y.float().permute(0, 3, 1, 2), yr.detach()) < 1e-2
    assert rel(dx.float().permute(0, 3, 1, 2), xr.grad) < 1e-2
    assert rel(dw - 0.5, wr.grad) < 2e-3


def test_inputs_of_a_1x1_stem(ops):
    """b2c_stem_nhwc64 / b2c_augment_nhwc64: the operand of ResNeXt's 1x1 stem conv (models/resnext.py:47) is the
    image itself in NHWC bf16, padded to 64 channels — from the fp32 NCHW batch a DataLoader yields, or from the uint8
    data set with the crop / flip / normalize of main.py:30-35 (same pixels as b2c_augment_nchw)."""
    from b200cifar import data
    n = 37
    g = torch.Generator(device="cuda").manual_seed(90)
    x = torch.randn(n, 3, 32, 32, device="cuda", generator=g)
    out = torch.full((n, 32, 32, 64), float("nan"), device="cuda", dtype=torch.bfloat16)
    ops.stem_im2col(x, out, taps=1)
    torch.cuda.synchronize()
    assert torch.equal(out[..., :3], x.permute(0, 2, 3, 1).to(torch.bfloat16)) and (out[..., 3:] == 0).all()
    rs = np.random.RandomState(5)
    images = torch.from_numpy(rs.randint(0, 256, size=(64, 32, 32, 3)).astype(np.uint8)).cuda()
    idx = torch.from_numpy(rs.randint(0, 64, size=(n,)).astype(np.int64)).cuda()
    ijf = torch.from_numpy(np.stack([rs.randint(0, 9, n), rs.randint(0, 9, n), rs.randint(0, 2, n)], 1).astype(np.int8)).cuda()
    ref = torch.empty(n, 3, 32, 32, device="cuda")
    ops.augment_nchw(images, idx, ijf, ref)
    out.fill_(float("nan"))
    ops.augment_im2col(images, idx, ijf, out, taps=1)
    torch.cuda.synchronize()
    assert torch.equal(out[..., :3], ref.permute(0, 2, 3, 1).to(torch.bfloat16)) and (out[..., 3:] == 0).all()


def test_head_over_a_padded_channel_layout(ops):
    """b2c_head_fwd_ld / b2c_head_bwd_ld: adaptive_avg_pool2d(1) -> Linear(368, 10) (models/regnet.py:104-106) over
    an activation stored with 384 channels per pixel; fp32 arithmetic, so tight against torch."""
    n, hw, C, ld, K = 37, 64, 368, 384, 10
    g = torch.Generator(device="cuda").manual_seed(95)
    act = torch.zeros(n, hw, ld, device="cuda", dtype=torch.bfloat16)
    act[..., :C] = torch.randn(n, hw, C, device="cuda", generator=g).to(torch.bfloat16)
    W = torch.randn(K, C, device="cuda", generator=g) * 0.05
    b = torch.randn(K, device="cuda", generator=g)
    pooled, logits = torch.empty(n, C, device="cuda"), torch.empty(n, K, device="cuda")
    ops.head_fwd(act, n, hw, C, W, b, K, pooled, logits, act_ld=ld)
    a32 = act[..., :C].float().requires_grad_()
    Wr, br = W.clone().requires_grad_(), b.clone().requires_grad_()
    ref = F.linear(a32.mean(1), Wr, br)
    dlog = torch.randn(n, K, device="cuda", generator=g)
    ref.backward(dlog)
    dact = torch.full((n, hw, ld), 7.0, device="cuda", dtype=torch.bfloat16)
    dW, db = torch.empty(K, C, device="cuda"), torch.empty(K, device="cuda")
    ops.head_bwd(dlog, pooled, W, n, hw, C, K, dact, dW, db, act_ld=ld)
    torch.cuda.synchronize()
    assert rel(logits, ref.detach()) < 1e-5 and rel(pooled, a32.detach().mean(1)) < 1e-6
    assert rel(dW, Wr.grad) < 1e-5 and rel(db, br.grad) < 1e-5
    assert rel(dact[..., :C].float(), a32.grad) < 5e-3 and (dact[..., C:] == 7.0).all()


@pytest.mark.parametrize("shape", [(9, 8, 128, 128, 8), (5, 32, 64, 64, 4), (130, 4, 512, 512, 32), (7, 8, 96, 128, 6)])
def test_squeeze_and_excitation(ops, shape):
    """b2c_se_fwd / b2c_se_bwd: w = sigmoid(fc2(relu(fc1(avg_pool(y))))); out = y * w + shortcut (models/senet.py:67-75)
    and its backward, against torch autograd in fp32 on the same bf16 inputs; the last case has 96 real channels in a
    128-channel layout."""
    n, hwside, C, ld, R = shape
    hw = hwside * hwside
    g = torch.Generator(device="cuda").manual_seed(97)
    y = torch.zeros(n, hw, ld, device="cuda", dtype=torch.bfloat16)
    y[..., :C] = torch.randn(n, hw, C, device="cuda", generator=g).to(torch.bfloat16)
    add = torch.zeros_like(y)
    add[..., :C] = torch.randn(n, hw, C, device="cuda", generator=g).to(torch.bfloat16)
    w1 = torch.randn(R, C, 1, 1, device="cuda", generator=g) * 0.2
    b1 = torch.randn(R, device="cuda", generator=g) * 0.1
    w2 = torch.randn(C, R, 1, 1, device="cuda", generator=g) * 0.5
    b2 = torch.randn(C, device="cuda", generator=g) * 0.1
    pooled, hidden, scale = torch.empty(n, ld, device="cuda"), torch.empty(n, R, device="cuda"), torch.empty(n, ld, device="cuda")
    out = torch.full_like(y, float("nan"))
    ops.se_fwd(y, n, hw, C, ld, R, w1, b1, w2, b2, pooled, hidden, scale, out, add=add)
    # torch
    yr = y[..., :C].float().permute(0, 2, 1).reshape(n, C, hwside, hwside).requires_grad_()
    leaves = [t.clone().requires_grad_() for t in (w1, b1, w2, b2)]
    w = F.avg_pool2d(yr, hwside)
    w = F.relu(F.conv2d(w, leaves[0], leaves[1]))
    w = torch.sigmoid(F.conv2d(w, leaves[2], leaves[3]))
    ref = yr * w + add[..., :C].float().permute(0, 2, 1).reshape(n, C, hwside, hwside)
    dout = torch.zeros_like(y)
    dout[..., :C] = torch.randn(n, hw, C, device="cuda", generator=g).to(torch.bfloat16)
    ref.backward(dout[..., :C].float().permute(0, 2, 1).reshape(n, C, hwside, hwside))
    ds, dz1, dp = torch.empty(n * ld, device="cuda"), torch.empty(n * R, device="cuda"), torch.empty(n * ld, device="cuda")
    grads = [torch.full_like(t, 0.25) for t in (w1, b1, w2, b2)]              # the parameter gradients accumulate
    dy = torch.full_like(y, float("nan"))
    ops.se_bwd(dout, y, n, hw, C, ld, R, w1, w2, pooled, hidden, scale, ds, dz1, dp, grads[0], grads[1], grads[2],
               grads[3], dy)
    torch.cuda.synchronize()
    to_nhwc = lambda t: t.reshape(n, C, hw).permute(0, 2, 1)
    assert rel(out[..., :C].float(), to_nhwc(ref.detach())) < 5e-3 and (out[..., C:] == 0).all()
    assert rel(scale[:, :C], w.detach().reshape(n, C)) < 1e-5 and (scale[:, C:] == 0).all()
    assert rel(dy[..., :C].float(), to_nhwc(yr.grad)) < 5e-3 and (dy[..., C:] == 0).all()
    for got, leaf in zip(grads, leaves):
        assert rel(got - 0.25, leaf.grad) < 1e-4, leaf.shape
