"""Functional fp32 restatement of the reference's CIFAR networks (test oracle; see oracle/__init__.py).

`build_resnet18_params(seed)` constructs torch.nn layers in exactly the order the reference's constructor
does (models/resnet.py:97-117 -> BasicBlock.__init__ :19-36), so the default initialisers consume the CPU
generator identically and the weights equal `torch.manual_seed(seed); models.resnet.ResNet18()` bit for bit.
`resnet_forward` restates ResNet.forward / BasicBlock.forward (models/resnet.py:47-51,132-140) with
functional ops; autograd of these fp32 ops is the backward oracle.
"""
from collections import OrderedDict

import torch
import torch.nn as nn
import torch.nn.functional as F

RESNET18_BLOCKS = (2, 2, 2, 2)
RESNET_PLANES = (64, 128, 256, 512)


def build_resnet_params(seed, num_blocks=RESNET18_BLOCKS, num_classes=10):
    """Returns (params, buffers): OrderedDicts keyed like the reference's state_dict."""
    torch.manual_seed(seed)
    params, buffers = OrderedDict(), OrderedDict()

    def conv(name, cin, cout, k, stride):
        m = nn.Conv2d(cin, cout, kernel_size=k, stride=stride, padding=k // 2, bias=False)
        params[name + ".weight"] = m.weight.detach().clone()

    def bn(name, c):
        m = nn.BatchNorm2d(c)
        params[name + ".weight"] = m.weight.detach().clone()
        params[name + ".bias"] = m.bias.detach().clone()
        buffers[name + ".running_mean"] = m.running_mean.clone()
        buffers[name + ".running_var"] = m.running_var.clone()
        buffers[name + ".num_batches_tracked"] = m.num_batches_tracked.clone()

    conv("conv1", 3, 64, 3, 1)                       # models/resnet.py:102-103
    bn("bn1", 64)                                    # :104
    in_planes = 64
    for li, (nb, planes) in enumerate(zip(num_blocks, RESNET_PLANES)):   # :105-108, _make_layer :111-117
        strides = [1 if li == 0 else 2] + [1] * (nb - 1)
        for bi, s in enumerate(strides):
            name = f"layer{li + 1}.{bi}"
            conv(name + ".conv1", in_planes, planes, 3, s)     # :23-24
            bn(name + ".bn1", planes)                          # :25
            conv(name + ".conv2", planes, planes, 3, 1)        # :26-27
            bn(name + ".bn2", planes)                          # :28
            if s != 1 or in_planes != planes:                  # :31
                conv(name + ".shortcut.0", in_planes, planes, 1, s)   # :33-34
                bn(name + ".shortcut.1", planes)                      # :35
            in_planes = planes
    lin = nn.Linear(in_planes, num_classes)          # :109
    params["linear.weight"] = lin.weight.detach().clone()
    params["linear.bias"] = lin.bias.detach().clone()
    return params, buffers


def _bn(x, name, params, buffers, training, momentum=0.1, eps=1e-5):
    rm, rv = buffers[name + ".running_mean"], buffers[name + ".running_var"]
    if training:
        buffers[name + ".num_batches_tracked"] += 1
    return F.batch_norm(x, rm, rv, params[name + ".weight"], params[name + ".bias"], training=training,
                        momentum=momentum, eps=eps)


def resnet_forward(x, params, buffers, training=True, num_blocks=RESNET18_BLOCKS, taps=None):
    """x: [N,3,32,32] fp32 NCHW. `taps` (optional dict) receives named intermediate tensors."""
    def tap(name, t):
        if taps is not None:
            taps[name] = t
        return t

    out = F.conv2d(x, params["conv1.weight"], padding=1)
    tap("conv1", out)
    out = tap("act0", F.relu(_bn(out, "bn1", params, buffers, training)))          # resnet.py:132
    in_planes = 64
    for li, (nb, planes) in enumerate(zip(num_blocks, RESNET_PLANES)):
        strides = [1 if li == 0 else 2] + [1] * (nb - 1)
        for bi, s in enumerate(strides):
            name = f"layer{li + 1}.{bi}"
            xin = out
            o = tap(name + ".conv1", F.conv2d(xin, params[name + ".conv1.weight"], stride=s, padding=1))
            o = tap(name + ".a1", F.relu(_bn(o, name + ".bn1", params, buffers, training)))   # :47
            o = tap(name + ".conv2", F.conv2d(o, params[name + ".conv2.weight"], padding=1))
            o = _bn(o, name + ".bn2", params, buffers, training)                   # :48
            if (name + ".shortcut.0.weight") in params:
                sc = tap(name + ".shortcut.0", F.conv2d(xin, params[name + ".shortcut.0.weight"], stride=s))
                sc = _bn(sc, name + ".shortcut.1", params, buffers, training)
            else:
                sc = xin
            out = tap(name + ".out", F.relu(o + sc))                               # :49-50
            in_planes = planes
    out = F.avg_pool2d(out, 4)                                                     # :137
    out = out.view(out.size(0), -1)                                                # :138
    return F.linear(out, params["linear.weight"], params["linear.bias"])          # :139


def train_step_grads(x, target, params, buffers, forward=resnet_forward, taps=None):
    """One forward + CrossEntropy + backward (main.py:101-104). Returns (loss, logits, grads by name)."""
    leaves = OrderedDict((k, v.detach().clone().requires_grad_(True)) for k, v in params.items())
    logits = forward(x, leaves, buffers, training=True, taps=taps)
    loss = F.cross_entropy(logits, target)                                         # main.py:86,103
    grads = torch.autograd.grad(loss, list(leaves.values()))
    return loss.detach(), logits.detach(), OrderedDict(zip(leaves.keys(), grads))


# ------------------------------------------------------------------------------------------------
# MobileNet v1 (models/mobilenet.py) — same primitives as MobileNetV2
# ------------------------------------------------------------------------------------------------
MOBILENET_CFG = [64, (128, 2), 128, (256, 2), 256, (512, 2), 512, 512, 512, 512, 512, (1024, 2), 1024]    # mobilenet.py:27


def _mobilenet_blocks():
    cin, out = 32, []
    for i, x in enumerate(MOBILENET_CFG):
        cout, s = (x, 1) if isinstance(x, int) else x
        out.append((f"layers.{i}", cin, cout, s))
        cin = cout
    return out


def build_mobilenet_params(seed):
    """Construction order of MobileNet.__init__ / Block.__init__ (mobilenet.py:13-18,31-34)."""
    torch.manual_seed(seed)
    params, buffers = OrderedDict(), OrderedDict()

    def conv(name, cin, cout, k, stride, groups=1):
        m = nn.Conv2d(cin, cout, kernel_size=k, stride=stride, padding=k // 2, groups=groups, bias=False)
        params[name + ".weight"] = m.weight.detach().clone()

    def bn(name, c):
        m = nn.BatchNorm2d(c)
        params[name + ".weight"] = m.weight.detach().clone()
        params[name + ".bias"] = m.bias.detach().clone()
        buffers[name + ".running_mean"] = m.running_mean.clone()
        buffers[name + ".running_var"] = m.running_var.clone()
        buffers[name + ".num_batches_tracked"] = m.num_batches_tracked.clone()
    conv("conv1", 3, 32, 3, 1); bn("bn1", 32)
    for name, cin, cout, s in _mobilenet_blocks():
        conv(name + ".conv1", cin, cin, 3, s, groups=cin); bn(name + ".bn1", cin)
        conv(name + ".conv2", cin, cout, 1, 1); bn(name + ".bn2", cout)
    lin = nn.Linear(1024, 10)
    params["linear.weight"] = lin.weight.detach().clone()
    params["linear.bias"] = lin.bias.detach().clone()
    return params, buffers


def mobilenet_forward(x, params, buffers, training=True, taps=None):
    out = F.conv2d(x, params["conv1.weight"], padding=1)
    if taps is not None:
        taps["conv1"] = out
    out = F.relu(_bn(out, "bn1", params, buffers, training))                                           # mobilenet.py:48
    for name, cin, cout, s in _mobilenet_blocks():
        out = F.conv2d(out, params[name + ".conv1.weight"], stride=s, padding=1, groups=cin)
        out = F.relu(_bn(out, name + ".bn1", params, buffers, training))                               # :21
        out = F.conv2d(out, params[name + ".conv2.weight"])
        if taps is not None:
            taps[name + ".conv2"] = out
        out = F.relu(_bn(out, name + ".bn2", params, buffers, training))                               # :22
    out = F.avg_pool2d(out, 2).view(out.size(0), -1)                                                   # :50-51
    return F.linear(out, params["linear.weight"], params["linear.bias"])


# ------------------------------------------------------------------------------------------------
# SimpleDLA (models/dla_simple.py) — the network main.py:71 constructs (first row)
# ------------------------------------------------------------------------------------------------
def _dla_tree(name, cin, cout, level, stride):
    if level == 1:
        left = ("block", name + ".left_tree", cin, cout, stride)
        right = ("block", name + ".right_tree", cout, cout, 1)
    else:
        left = _dla_tree(name + ".left_tree", cin, cout, level - 1, stride)
        right = _dla_tree(name + ".right_tree", cout, cout, level - 1, 1)
    return ("tree", name, cout, left, right)


DLA_TREES = [_dla_tree("layer3", 32, 64, 1, 1), _dla_tree("layer4", 64, 128, 2, 2),
             _dla_tree("layer5", 128, 256, 2, 2), _dla_tree("layer6", 256, 512, 1, 2)]      # dla_simple.py:98-101


def build_simpledla_params(seed):
    """Construction order of SimpleDLA.__init__ (dla_simple.py:80-102): base, layer1, layer2, then per Tree the Root
    first, then left_tree, right_tree (Tree.__init__ :61-70), each BasicBlock as in :19-36; Linear last."""
    torch.manual_seed(seed)
    params, buffers = OrderedDict(), OrderedDict()

    def conv(name, cin, cout, k, stride):
        m = nn.Conv2d(cin, cout, kernel_size=k, stride=stride, padding=k // 2, bias=False)
        params[name + ".weight"] = m.weight.detach().clone()

    def bn(name, c):
        m = nn.BatchNorm2d(c)
        params[name + ".weight"] = m.weight.detach().clone()
        params[name + ".bias"] = m.bias.detach().clone()
        buffers[name + ".running_mean"] = m.running_mean.clone()
        buffers[name + ".running_var"] = m.running_var.clone()
        buffers[name + ".num_batches_tracked"] = m.num_batches_tracked.clone()

    def node(n):
        if n[0] == "block":
            _, name, cin, cout, s = n
            conv(name + ".conv1", cin, cout, 3, s); bn(name + ".bn1", cout)
            conv(name + ".conv2", cout, cout, 3, 1); bn(name + ".bn2", cout)
            if s != 1 or cin != cout:
                conv(name + ".shortcut.0", cin, cout, 1, s); bn(name + ".shortcut.1", cout)
        else:
            _, name, cout, left, right = n
            conv(name + ".root.conv", 2 * cout, cout, 1, 1); bn(name + ".root.bn", cout)
            node(left); node(right)
    for name, cin, cout in (("base", 3, 16), ("layer1", 16, 16), ("layer2", 16, 32)):
        conv(name + ".0", cin, cout, 3, 1); bn(name + ".1", cout)
    for t in DLA_TREES:
        node(t)
    lin = nn.Linear(512, 10)
    params["linear.weight"] = lin.weight.detach().clone()
    params["linear.bias"] = lin.bias.detach().clone()
    return params, buffers


def simpledla_forward(x, params, buffers, training=True, taps=None):
    def block(n, xin):
        _, name, cin, cout, s = n
        o = F.relu(_bn(F.conv2d(xin, params[name + ".conv1.weight"], stride=s, padding=1), name + ".bn1", params,
                       buffers, training))                                                 # dla_simple.py:37
        o = _bn(F.conv2d(o, params[name + ".conv2.weight"], padding=1), name + ".bn2", params, buffers, training)
        if (name + ".shortcut.0.weight") in params:
            sc = _bn(F.conv2d(xin, params[name + ".shortcut.0.weight"], stride=s), name + ".shortcut.1", params, buffers,
                     training)
        else:
            sc = xin
        return F.relu(o + sc)                                                              # :39-40

    def node(n, xin):
        if n[0] == "block":
            return block(n, xin)
        _, name, cout, left, right = n
        out1 = node(left, xin)                                                             # :73
        out2 = node(right, out1)                                                           # :74
        y = F.conv2d(torch.cat([out1, out2], 1), params[name + ".root.conv.weight"])       # :52-53
        if taps is not None:
            taps[name + ".root.conv"] = y
        return F.relu(_bn(y, name + ".root.bn", params, buffers, training))                # :54
    out = x
    for name in ("base", "layer1", "layer2"):                                              # :108-110
        out = F.conv2d(out, params[name + ".0.weight"], padding=1)
        if taps is not None:
            taps[name + ".0"] = out
        out = F.relu(_bn(out, name + ".1", params, buffers, training))
    for t in DLA_TREES:                                                                    # :111-114
        out = node(t, out)
    out = F.avg_pool2d(out, 4).view(out.size(0), -1)                                       # :115-116
    return F.linear(out, params["linear.weight"], params["linear.bias"])


# ------------------------------------------------------------------------------------------------
# VGG (models/vgg.py) — BASELINE config 3
# ------------------------------------------------------------------------------------------------
VGG_CFG = {
    "VGG11": [64, "M", 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
    "VGG13": [64, 64, "M", 128, 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
    "VGG16": [64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512, "M"],
    "VGG19": [64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M", 512, 512, 512, 512, "M"],
}


def build_vgg_params(seed, name="VGG16"):
    """Same construction order as VGG.__init__ (models/vgg.py:15-18): features first, then the classifier."""
    torch.manual_seed(seed)
    params, buffers = OrderedDict(), OrderedDict()
    i, cin = 0, 3
    for x in VGG_CFG[name]:                                        # _make_layers, vgg.py:26-38
        if x == "M":
            i += 1
            continue
        conv = nn.Conv2d(cin, x, kernel_size=3, padding=1)         # bias=True (vgg.py:33)
        bn = nn.BatchNorm2d(x)
        params[f"features.{i}.weight"] = conv.weight.detach().clone()
        params[f"features.{i}.bias"] = conv.bias.detach().clone()
        params[f"features.{i + 1}.weight"] = bn.weight.detach().clone()
        params[f"features.{i + 1}.bias"] = bn.bias.detach().clone()
        buffers[f"features.{i + 1}.running_mean"] = bn.running_mean.clone()
        buffers[f"features.{i + 1}.running_var"] = bn.running_var.clone()
        buffers[f"features.{i + 1}.num_batches_tracked"] = bn.num_batches_tracked.clone()
        cin = x
        i += 3
    lin = nn.Linear(512, 10)                                       # vgg.py:18
    params["classifier.weight"] = lin.weight.detach().clone()
    params["classifier.bias"] = lin.bias.detach().clone()
    return params, buffers


def vgg_forward(x, params, buffers, training=True, name="VGG16", taps=None):
    out, i = x, 0
    for c in VGG_CFG[name]:
        if c == "M":
            out = F.max_pool2d(out, kernel_size=2, stride=2)       # vgg.py:31
            i += 1
            continue
        out = F.conv2d(out, params[f"features.{i}.weight"], params[f"features.{i}.bias"], padding=1)
        if taps is not None:
            taps[f"features.{i}"] = out
        out = F.relu(_bn(out, f"features.{i + 1}", params, buffers, training))     # vgg.py:33-35
        i += 3
    out = out.view(out.size(0), -1)                                # AvgPool2d(1,1) is the identity; vgg.py:22
    return F.linear(out, params["classifier.weight"], params["classifier.bias"])


# ------------------------------------------------------------------------------------------------
# PreActResNet-18 (models/preact_resnet.py) — BASELINE config 4
# ------------------------------------------------------------------------------------------------
def build_preact_params(seed, num_blocks=RESNET18_BLOCKS):
    """Construction order of PreActResNet.__init__ / PreActBlock.__init__ (preact_resnet.py:16-26,68-75)."""
    torch.manual_seed(seed)
    params, buffers = OrderedDict(), OrderedDict()

    def conv(name, cin, cout, k, stride):
        m = nn.Conv2d(cin, cout, kernel_size=k, stride=stride, padding=k // 2, bias=False)
        params[name + ".weight"] = m.weight.detach().clone()

    def bn(name, c):
        m = nn.BatchNorm2d(c)
        params[name + ".weight"] = m.weight.detach().clone()
        params[name + ".bias"] = m.bias.detach().clone()
        buffers[name + ".running_mean"] = m.running_mean.clone()
        buffers[name + ".running_var"] = m.running_var.clone()
        buffers[name + ".num_batches_tracked"] = m.num_batches_tracked.clone()

    conv("conv1", 3, 64, 3, 1)
    in_planes = 64
    for li, (nb, planes) in enumerate(zip(num_blocks, RESNET_PLANES)):
        strides = [1 if li == 0 else 2] + [1] * (nb - 1)
        for bi, s in enumerate(strides):
            name = f"layer{li + 1}.{bi}"
            bn(name + ".bn1", in_planes)                           # :18
            conv(name + ".conv1", in_planes, planes, 3, s)         # :19
            bn(name + ".bn2", planes)                              # :20
            conv(name + ".conv2", planes, planes, 3, 1)            # :21
            if s != 1 or in_planes != planes:                      # :23-26
                conv(name + ".shortcut.0", in_planes, planes, 1, s)
            in_planes = planes
    lin = nn.Linear(in_planes, 10)
    params["linear.weight"] = lin.weight.detach().clone()
    params["linear.bias"] = lin.bias.detach().clone()
    return params, buffers


def preact_forward(x, params, buffers, training=True, num_blocks=RESNET18_BLOCKS, taps=None):
    out = F.conv2d(x, params["conv1.weight"], padding=1)           # preact_resnet.py:86
    if taps is not None:
        taps["conv1"] = out
    for li, (nb, planes) in enumerate(zip(num_blocks, RESNET_PLANES)):
        strides = [1 if li == 0 else 2] + [1] * (nb - 1)
        for bi, s in enumerate(strides):
            name = f"layer{li + 1}.{bi}"
            xin = out
            o = F.relu(_bn(xin, name + ".bn1", params, buffers, training))          # :29
            if (name + ".shortcut.0.weight") in params:
                sc = F.conv2d(o, params[name + ".shortcut.0.weight"], stride=s)     # :30
            else:
                sc = xin
            o = F.conv2d(o, params[name + ".conv1.weight"], stride=s, padding=1)    # :31
            if taps is not None:
                taps[name + ".conv1"] = o
            o = F.conv2d(F.relu(_bn(o, name + ".bn2", params, buffers, training)),
                         params[name + ".conv2.weight"], padding=1)                 # :32
            out = o + sc                                                            # :33
            if taps is not None:
                taps[name + ".out"] = out
    out = F.avg_pool2d(out, 4)                                                      # :91
    out = out.view(out.size(0), -1)
    return F.linear(out, params["linear.weight"], params["linear.bias"])


# ------------------------------------------------------------------------------------------------
# MobileNetV2 (models/mobilenetv2.py) — BASELINE config 5
# ------------------------------------------------------------------------------------------------
MOBILENETV2_CFG = [(1, 16, 1, 1), (6, 24, 2, 1), (6, 32, 3, 2), (6, 64, 4, 2), (6, 96, 3, 1), (6, 160, 3, 2),
                   (6, 320, 1, 1)]        # mobilenetv2.py:42-48


def _mbv2_blocks():
    cin, i = 32, 0
    for expansion, cout, nb, stride in MOBILENETV2_CFG:
        for s in [stride] + [1] * (nb - 1):
            yield f"layers.{i}", cin, cout, expansion * cin, s
            cin = cout
            i += 1


def build_mobilenetv2_params(seed):
    """Construction order of MobileNetV2.__init__ / Block.__init__ (mobilenetv2.py:13-30,50-58)."""
    torch.manual_seed(seed)
    params, buffers = OrderedDict(), OrderedDict()

    def conv(name, cin, cout, k, stride, groups=1):
        m = nn.Conv2d(cin, cout, kernel_size=k, stride=stride, padding=k // 2, groups=groups, bias=False)
        params[name + ".weight"] = m.weight.detach().clone()

    def bn(name, c):
        m = nn.BatchNorm2d(c)
        params[name + ".weight"] = m.weight.detach().clone()
        params[name + ".bias"] = m.bias.detach().clone()
        buffers[name + ".running_mean"] = m.running_mean.clone()
        buffers[name + ".running_var"] = m.running_var.clone()
        buffers[name + ".num_batches_tracked"] = m.num_batches_tracked.clone()

    conv("conv1", 3, 32, 3, 1)                                   # :52
    bn("bn1", 32)
    for name, cin, cout, planes, s in _mbv2_blocks():
        conv(name + ".conv1", cin, planes, 1, 1)                 # :18
        bn(name + ".bn1", planes)
        conv(name + ".conv2", planes, planes, 3, s, groups=planes)   # :20 depthwise
        bn(name + ".bn2", planes)
        conv(name + ".conv3", planes, cout, 1, 1)                # :22
        bn(name + ".bn3", cout)
        if s == 1 and cin != cout:                               # :26-30
            conv(name + ".shortcut.0", cin, cout, 1, 1)
            bn(name + ".shortcut.1", cout)
    conv("conv2", 320, 1280, 1, 1)                               # :55
    bn("bn2", 1280)
    lin = nn.Linear(1280, 10)
    params["linear.weight"] = lin.weight.detach().clone()
    params["linear.bias"] = lin.bias.detach().clone()
    return params, buffers


def mobilenetv2_forward(x, params, buffers, training=True, taps=None):
    out = F.conv2d(x, params["conv1.weight"], padding=1)
    if taps is not None:
        taps["conv1"] = out
    out = F.relu(_bn(out, "bn1", params, buffers, training))                                   # :70
    for name, cin, cout, planes, s in _mbv2_blocks():
        xin = out
        o = F.conv2d(xin, params[name + ".conv1.weight"])
        if taps is not None:
            taps[name + ".conv1"] = o
        o = F.relu(_bn(o, name + ".bn1", params, buffers, training))                           # :33
        o = F.conv2d(o, params[name + ".conv2.weight"], stride=s, padding=1, groups=planes)
        if taps is not None:
            taps[name + ".conv2"] = o
        o = F.relu(_bn(o, name + ".bn2", params, buffers, training))                           # :34
        o = _bn(F.conv2d(o, params[name + ".conv3.weight"]), name + ".bn3", params, buffers, training)   # :35
        if s == 1:                                                                             # :36
            if (name + ".shortcut.0.weight") in params:
                sc = _bn(F.conv2d(xin, params[name + ".shortcut.0.weight"]), name + ".shortcut.1", params, buffers,
                         training)
            else:
                sc = xin
            o = o + sc
        out = o
    out = F.relu(_bn(F.conv2d(out, params["conv2.weight"]), "bn2", params, buffers, training)) # :72
    out = F.avg_pool2d(out, 4).view(out.size(0), -1)                                           # :74-75
    return F.linear(out, params["linear.weight"], params["linear.bias"])


# ------------------------------------------------------------------------------------------------
# Bottleneck ResNets (models/resnet.py:54-93,150-160) — ResNet-50/101/152; main_dist.py:136 builds ResNet152
# ------------------------------------------------------------------------------------------------
RESNET50_BLOCKS = (3, 4, 6, 3)


def _bottleneck_blocks(num_blocks):
    cin = 64
    for li, (nb, planes) in enumerate(zip(num_blocks, RESNET_PLANES)):
        for bi, s in enumerate([1 if li == 0 else 2] + [1] * (nb - 1)):
            yield f"layer{li + 1}.{bi}", cin, planes, 4 * planes, s
            cin = 4 * planes


def build_bottleneck_params(seed, num_blocks=RESNET50_BLOCKS):
    torch.manual_seed(seed)
    params, buffers = OrderedDict(), OrderedDict()

    def conv(name, cin, cout, k, stride):
        m = nn.Conv2d(cin, cout, kernel_size=k, stride=stride, padding=k // 2, bias=False)
        params[name + ".weight"] = m.weight.detach().clone()

    def bn(name, c):
        m = nn.BatchNorm2d(c)
        params[name + ".weight"] = m.weight.detach().clone()
        params[name + ".bias"] = m.bias.detach().clone()
        buffers[name + ".running_mean"] = m.running_mean.clone()
        buffers[name + ".running_var"] = m.running_var.clone()
        buffers[name + ".num_batches_tracked"] = m.num_batches_tracked.clone()

    conv("conv1", 3, 64, 3, 1)
    bn("bn1", 64)
    cin_final = 64
    for name, cin, planes, cout, s in _bottleneck_blocks(num_blocks):
        conv(name + ".conv1", cin, planes, 1, 1)          # resnet.py:61
        bn(name + ".bn1", planes)
        conv(name + ".conv2", planes, planes, 3, s)       # :63-64
        bn(name + ".bn2", planes)
        conv(name + ".conv3", planes, cout, 1, 1)         # :66-67
        bn(name + ".bn3", cout)
        if s != 1 or cin != cout:                         # :71-76
            conv(name + ".shortcut.0", cin, cout, 1, s)
            bn(name + ".shortcut.1", cout)
        cin_final = cout
    lin = nn.Linear(cin_final, 10)
    params["linear.weight"] = lin.weight.detach().clone()
    params["linear.bias"] = lin.bias.detach().clone()
    return params, buffers


def bottleneck_forward(x, params, buffers, training=True, num_blocks=RESNET50_BLOCKS, taps=None):
    out = F.conv2d(x, params["conv1.weight"], padding=1)
    if taps is not None:
        taps["conv1"] = out
    out = F.relu(_bn(out, "bn1", params, buffers, training))
    for name, cin, planes, cout, s in _bottleneck_blocks(num_blocks):
        xin = out
        o = F.conv2d(xin, params[name + ".conv1.weight"])
        if taps is not None:
            taps[name + ".conv1"] = o
        o = F.relu(_bn(o, name + ".bn1", params, buffers, training))                                  # :87
        o = F.relu(_bn(F.conv2d(o, params[name + ".conv2.weight"], stride=s, padding=1), name + ".bn2", params,
                       buffers, training))                                                            # :88
        o = _bn(F.conv2d(o, params[name + ".conv3.weight"]), name + ".bn3", params, buffers, training)   # :89
        if (name + ".shortcut.0.weight") in params:
            sc = _bn(F.conv2d(xin, params[name + ".shortcut.0.weight"], stride=s), name + ".shortcut.1", params, buffers,
                     training)
        else:
            sc = xin
        out = F.relu(o + sc)                                                                          # :90-91
    out = F.avg_pool2d(out, 4).view(out.size(0), -1)
    return F.linear(out, params["linear.weight"], params["linear.bias"])


# ------------------------------------------------------------------------------------------------
# ResNeXt-29 (models/resnext.py:10-86): grouped 3x3 convs, 1x1 stem, three stages, avg_pool2d(8)
# ------------------------------------------------------------------------------------------------
RESNEXT29_BLOCKS = (3, 3, 3)


def _resnext_blocks(num_blocks, cardinality, bottleneck_width):
    cin, bw = 64, bottleneck_width
    for li, nb in enumerate(num_blocks):                      # resnext.py:57-66: the width doubles after each stage
        for bi, s in enumerate([1 if li == 0 else 2] + [1] * (nb - 1)):
            gw = cardinality * bw
            yield f"layer{li + 1}.{bi}", cin, gw, 2 * gw, s
            cin = 2 * gw
        bw *= 2


def build_resnext_params(seed, cardinality=2, bottleneck_width=64, num_blocks=RESNEXT29_BLOCKS):
    """Seeded parameters equal to `torch.manual_seed(seed); models.resnext.ResNeXt(num_blocks, cardinality,
    bottleneck_width)` (resnext.py:14-28,42-54): layers are constructed in the reference's order."""
    torch.manual_seed(seed)
    params, buffers = OrderedDict(), OrderedDict()

    def conv(name, cin, cout, k, stride, groups=1):
        m = nn.Conv2d(cin, cout, kernel_size=k, stride=stride, padding=k // 2, groups=groups, bias=False)
        params[name + ".weight"] = m.weight.detach().clone()

    def bn(name, c):
        m = nn.BatchNorm2d(c)
        params[name + ".weight"] = m.weight.detach().clone()
        params[name + ".bias"] = m.bias.detach().clone()
        buffers[name + ".running_mean"] = m.running_mean.clone()
        buffers[name + ".running_var"] = m.running_var.clone()
        buffers[name + ".num_batches_tracked"] = m.num_batches_tracked.clone()

    conv("conv1", 3, 64, 1, 1)                                # resnext.py:47: kernel_size=1
    bn("bn1", 64)
    cin_final = 64
    for name, cin, gw, cout, s in _resnext_blocks(num_blocks, cardinality, bottleneck_width):
        conv(name + ".conv1", cin, gw, 1, 1)                  # :17
        bn(name + ".bn1", gw)
        conv(name + ".conv2", gw, gw, 3, s, groups=cardinality)   # :19
        bn(name + ".bn2", gw)
        conv(name + ".conv3", gw, cout, 1, 1)                 # :21
        bn(name + ".bn3", cout)
        if s != 1 or cin != cout:                             # :24-29
            conv(name + ".shortcut.0", cin, cout, 1, s)
            bn(name + ".shortcut.1", cout)
        cin_final = cout
    lin = nn.Linear(cin_final, 10)                            # :54: cardinality * bottleneck_width * 8
    params["linear.weight"] = lin.weight.detach().clone()
    params["linear.bias"] = lin.bias.detach().clone()
    return params, buffers


def resnext_forward(x, params, buffers, training=True, cardinality=2, bottleneck_width=64,
                    num_blocks=RESNEXT29_BLOCKS, taps=None):
    out = F.conv2d(x, params["conv1.weight"])                 # resnext.py:69
    if taps is not None:
        taps["conv1"] = out
    out = F.relu(_bn(out, "bn1", params, buffers, training))
    for name, cin, gw, cout, s in _resnext_blocks(num_blocks, cardinality, bottleneck_width):
        xin = out
        o = F.conv2d(xin, params[name + ".conv1.weight"])
        if taps is not None:
            taps[name + ".conv1"] = o
        o = F.relu(_bn(o, name + ".bn1", params, buffers, training))                                      # :32
        o = F.conv2d(o, params[name + ".conv2.weight"], stride=s, padding=1, groups=cardinality)
        if taps is not None:
            taps[name + ".conv2"] = o
        o = F.relu(_bn(o, name + ".bn2", params, buffers, training))                                      # :33
        o = _bn(F.conv2d(o, params[name + ".conv3.weight"]), name + ".bn3", params, buffers, training)   # :34
        if (name + ".shortcut.0.weight") in params:
            sc = _bn(F.conv2d(xin, params[name + ".shortcut.0.weight"], stride=s), name + ".shortcut.1", params, buffers,
                     training)
        else:
            sc = xin
        out = F.relu(o + sc)                                                                              # :35-36
    out = F.avg_pool2d(out, 8).view(out.size(0), -1)          # :74-75
    return F.linear(out, params["linear.weight"], params["linear.bias"])


def resnext_32x4d_forward(x, params, buffers, training=True, taps=None):
    return resnext_forward(x, params, buffers, training, cardinality=32, bottleneck_width=4, taps=taps)


def resnext_4x64d_forward(x, params, buffers, training=True, taps=None):
    return resnext_forward(x, params, buffers, training, cardinality=4, bottleneck_width=64, taps=taps)


def build_resnext_4x64d_params(seed):
    return build_resnext_params(seed, cardinality=4, bottleneck_width=64)


def build_resnext_32x4d_params(seed):
    return build_resnext_params(seed, cardinality=32, bottleneck_width=4)


# ------------------------------------------------------------------------------------------------
# RegNetX (models/regnet.py:29-134, se_ratio 0): the same block with widths off the 64-channel grid
# ------------------------------------------------------------------------------------------------
REGNETX_CFG = {
    "RegNetX_200MF": dict(depths=(1, 1, 4, 7), widths=(24, 56, 152, 368), strides=(1, 1, 2, 2), group_width=8),   # :113-122
    "RegNetX_400MF": dict(depths=(1, 2, 7, 12), widths=(32, 64, 160, 384), strides=(1, 1, 2, 2), group_width=16),  # :125-134
    "RegNetY_400MF": dict(depths=(1, 2, 7, 12), widths=(32, 64, 160, 384), strides=(1, 1, 2, 2), group_width=16,
                          se_ratio=0.25),                                                                           # :137-146
}


def _regnet_blocks(cfg):
    cin = 64
    for li, (depth, width, stride) in enumerate(zip(cfg["depths"], cfg["widths"], cfg["strides"])):   # regnet.py:82-96
        for bi in range(depth):
            yield f"layer{li + 1}.{bi}", cin, width, (stride if bi == 0 else 1), width // cfg["group_width"]
            cin = width


def build_regnetx_params(seed, name="RegNetX_200MF"):
    """Seeded parameters equal to `torch.manual_seed(seed); models.regnet.<name>()` (regnet.py:30-54,67-79)."""
    cfg = REGNETX_CFG[name]
    torch.manual_seed(seed)
    params, buffers = OrderedDict(), OrderedDict()

    def conv(cname, cin, cout, k, stride, groups=1, bias=False):
        m = nn.Conv2d(cin, cout, kernel_size=k, stride=stride, padding=k // 2, groups=groups, bias=bias)
        params[cname + ".weight"] = m.weight.detach().clone()
        if bias:
            params[cname + ".bias"] = m.bias.detach().clone()

    def bn(bname, c):
        m = nn.BatchNorm2d(c)
        params[bname + ".weight"] = m.weight.detach().clone()
        params[bname + ".bias"] = m.bias.detach().clone()
        buffers[bname + ".running_mean"] = m.running_mean.clone()
        buffers[bname + ".running_var"] = m.running_var.clone()
        buffers[bname + ".num_batches_tracked"] = m.num_batches_tracked.clone()

    conv("conv1", 3, 64, 3, 1)
    bn("bn1", 64)
    cin_final = 64
    for bname, cin, width, s, g in _regnet_blocks(cfg):
        conv(bname + ".conv1", cin, width, 1, 1)               # :34 (w_b = w_out: bottleneck_ratio 1)
        bn(bname + ".bn1", width)
        conv(bname + ".conv2", width, width, 3, s, groups=g)   # :37-39
        bn(bname + ".bn2", width)
        if cfg.get("se_ratio", 0) > 0:                          # :42-45, SE.__init__ :15-18
            w_se = int(round(cin * cfg["se_ratio"]))
            conv(bname + ".se.se1", width, w_se, 1, 1, bias=True)
            conv(bname + ".se.se2", w_se, width, 1, 1, bias=True)
        conv(bname + ".conv3", width, width, 1, 1)             # :46
        bn(bname + ".bn3", width)
        if s != 1 or cin != width:                             # :49-55
            conv(bname + ".shortcut.0", cin, width, 1, s)
            bn(bname + ".shortcut.1", width)
        cin_final = width
    lin = nn.Linear(cin_final, 10)
    params["linear.weight"] = lin.weight.detach().clone()
    params["linear.bias"] = lin.bias.detach().clone()
    return params, buffers


def regnetx_forward(x, params, buffers, training=True, name="RegNetX_200MF", taps=None):
    cfg = REGNETX_CFG[name]
    out = F.conv2d(x, params["conv1.weight"], padding=1)
    if taps is not None:
        taps["conv1"] = out
    out = F.relu(_bn(out, "bn1", params, buffers, training))                                             # regnet.py:99
    for bname, cin, width, s, g in _regnet_blocks(cfg):
        xin = out
        o = F.conv2d(xin, params[bname + ".conv1.weight"])
        if taps is not None:
            taps[bname + ".conv1"] = o
        o = F.relu(_bn(o, bname + ".bn1", params, buffers, training))                                     # :58
        o = F.conv2d(o, params[bname + ".conv2.weight"], stride=s, padding=1, groups=g)
        if taps is not None:
            taps[bname + ".conv2"] = o
        o = F.relu(_bn(o, bname + ".bn2", params, buffers, training))                                     # :59
        if (bname + ".se.se1.weight") in params:                                                          # :60-61, :20-25
            w = F.adaptive_avg_pool2d(o, (1, 1))
            w = F.relu(F.conv2d(w, params[bname + ".se.se1.weight"], params[bname + ".se.se1.bias"]))
            w = torch.sigmoid(F.conv2d(w, params[bname + ".se.se2.weight"], params[bname + ".se.se2.bias"]))
            o = o * w
        o = _bn(F.conv2d(o, params[bname + ".conv3.weight"]), bname + ".bn3", params, buffers, training)  # :62
        if (bname + ".shortcut.0.weight") in params:
            sc = _bn(F.conv2d(xin, params[bname + ".shortcut.0.weight"], stride=s), bname + ".shortcut.1", params,
                     buffers, training)
        else:
            sc = xin
        out = F.relu(o + sc)                                                                              # :63-64
    out = F.adaptive_avg_pool2d(out, (1, 1)).view(out.size(0), -1)                                        # :104-105
    return F.linear(out, params["linear.weight"], params["linear.bias"])


def build_regnetx_400mf_params(seed):
    return build_regnetx_params(seed, "RegNetX_400MF")


def regnetx_400mf_forward(x, params, buffers, training=True, taps=None):
    return regnetx_forward(x, params, buffers, training, "RegNetX_400MF", taps)


# ------------------------------------------------------------------------------------------------
# DLA (models/dla.py): level-2 trees aggregate FOUR tensors in their Root
# ------------------------------------------------------------------------------------------------
DLA_FULL_TREES = [("layer3", 32, 64, 1, 1), ("layer4", 64, 128, 2, 2), ("layer5", 128, 256, 2, 2),
                  ("layer6", 256, 512, 1, 2)]          # (name, in, out, level, stride), dla.py:108-111


def build_dla_params(seed):
    """Construction order of DLA.__init__ (dla.py:88-112): base, layer1, layer2, then per Tree (:54-69) the Root
    first; level 1: left_node, right_node; level 2: level_1 (a level-1 Tree), prev_root, left_node, right_node."""
    torch.manual_seed(seed)
    params, buffers = OrderedDict(), OrderedDict()

    def conv(name, cin, cout, k, stride):
        m = nn.Conv2d(cin, cout, kernel_size=k, stride=stride, padding=k // 2, bias=False)
        params[name + ".weight"] = m.weight.detach().clone()

    def bn(name, c):
        m = nn.BatchNorm2d(c)
        params[name + ".weight"] = m.weight.detach().clone()
        params[name + ".bias"] = m.bias.detach().clone()
        buffers[name + ".running_mean"] = m.running_mean.clone()
        buffers[name + ".running_var"] = m.running_var.clone()
        buffers[name + ".num_batches_tracked"] = m.num_batches_tracked.clone()

    def block(name, cin, cout, s):                            # dla.py:14-30
        conv(name + ".conv1", cin, cout, 3, s); bn(name + ".bn1", cout)
        conv(name + ".conv2", cout, cout, 3, 1); bn(name + ".bn2", cout)
        if s != 1 or cin != cout:
            conv(name + ".shortcut.0", cin, cout, 1, s); bn(name + ".shortcut.1", cout)

    def tree(name, cin, cout, level, s):
        if level == 1:
            conv(name + ".root.conv", 2 * cout, cout, 1, 1); bn(name + ".root.bn", cout)          # :58
            block(name + ".left_node", cin, cout, s)
            block(name + ".right_node", cout, cout, 1)
        else:
            conv(name + ".root.conv", (level + 2) * cout, cout, 1, 1); bn(name + ".root.bn", cout)   # :62
            for i in reversed(range(1, level)):                                                    # :63-66
                tree(name + f".level_{i}", cin, cout, i, s)
            block(name + ".prev_root", cin, cout, s)                                               # :67
            block(name + ".left_node", cout, cout, 1)
            block(name + ".right_node", cout, cout, 1)
    for name, cin, cout in (("base", 3, 16), ("layer1", 16, 16), ("layer2", 16, 32)):
        conv(name + ".0", cin, cout, 3, 1); bn(name + ".1", cout)
    for name, cin, cout, level, s in DLA_FULL_TREES:
        tree(name, cin, cout, level, s)
    lin = nn.Linear(512, 10)
    params["linear.weight"] = lin.weight.detach().clone()
    params["linear.bias"] = lin.bias.detach().clone()
    return params, buffers


def dla_forward(x, params, buffers, training=True, taps=None):
    def block(name, xin, s):
        o = F.relu(_bn(F.conv2d(xin, params[name + ".conv1.weight"], stride=s, padding=1), name + ".bn1", params,
                       buffers, training))                                                 # dla.py:33
        o = _bn(F.conv2d(o, params[name + ".conv2.weight"], padding=1), name + ".bn2", params, buffers, training)
        if (name + ".shortcut.0.weight") in params:
            sc = _bn(F.conv2d(xin, params[name + ".shortcut.0.weight"], stride=s), name + ".shortcut.1", params, buffers,
                     training)
        else:
            sc = xin
        return F.relu(o + sc)                                                              # :35-36

    def tree(name, xin, level, s):
        xs = [block(name + ".prev_root", xin, s)] if level > 1 else []                     # :72
        h = xin
        for i in reversed(range(1, level)):                                                # :73-76
            h = tree(name + f".level_{i}", h, i, s)
            xs.append(h)
        h = block(name + ".left_node", h, s if level == 1 else 1)                          # :77-78
        xs.append(h)
        h = block(name + ".right_node", h, 1)                                              # :79-80
        xs.append(h)
        y = F.conv2d(torch.cat(xs, 1), params[name + ".root.conv.weight"])                 # :48-49
        if taps is not None:
            taps[name + ".root.conv"] = y
        return F.relu(_bn(y, name + ".root.bn", params, buffers, training))                # :50
    out = x
    for name in ("base", "layer1", "layer2"):                                              # :115-117
        out = F.conv2d(out, params[name + ".0.weight"], padding=1)
        if taps is not None:
            taps[name + ".0"] = out
        out = F.relu(_bn(out, name + ".1", params, buffers, training))
    for name, cin, cout, level, s in DLA_FULL_TREES:                                       # :118-121
        out = tree(name, out, level, s)
    out = F.avg_pool2d(out, 4).view(out.size(0), -1)                                       # :122-123
    return F.linear(out, params["linear.weight"], params["linear.bias"])


# ------------------------------------------------------------------------------------------------
# SENet18 (models/senet.py): pre-activation blocks with squeeze-and-excitation
# ------------------------------------------------------------------------------------------------
def _senet_blocks(num_blocks=RESNET18_BLOCKS):
    cin = 64
    for li, (nb, pl) in enumerate(zip(num_blocks, RESNET_PLANES)):
        for bi, s in enumerate([1 if li == 0 else 2] + [1] * (nb - 1)):
            yield f"layer{li + 1}.{bi}", cin, pl, s
            cin = pl


def build_senet_params(seed, num_blocks=RESNET18_BLOCKS):
    """Seeded parameters equal to `torch.manual_seed(seed); models.senet.SENet18()`: SENet.__init__ (senet.py:82-94)
    with PreActBlock.__init__ (:45-60) — bn1, conv1, bn2, conv2, shortcut conv, fc1, fc2 (1x1 convs with bias)."""
    torch.manual_seed(seed)
    params, buffers = OrderedDict(), OrderedDict()

    def conv(name, cin, cout, k, stride, bias=False):
        m = nn.Conv2d(cin, cout, kernel_size=k, stride=stride, padding=k // 2, bias=bias)
        params[name + ".weight"] = m.weight.detach().clone()
        if bias:
            params[name + ".bias"] = m.bias.detach().clone()

    def bn(name, c):
        m = nn.BatchNorm2d(c)
        params[name + ".weight"] = m.weight.detach().clone()
        params[name + ".bias"] = m.bias.detach().clone()
        buffers[name + ".running_mean"] = m.running_mean.clone()
        buffers[name + ".running_var"] = m.running_var.clone()
        buffers[name + ".num_batches_tracked"] = m.num_batches_tracked.clone()

    conv("conv1", 3, 64, 3, 1)
    bn("bn1", 64)
    for name, cin, pl, s in _senet_blocks(num_blocks):
        bn(name + ".bn1", cin)
        conv(name + ".conv1", cin, pl, 3, s)
        bn(name + ".bn2", pl)
        conv(name + ".conv2", pl, pl, 3, 1)
        if s != 1 or cin != pl:
            conv(name + ".shortcut.0", cin, pl, 1, s)
        conv(name + ".fc1", pl, pl // 16, 1, 1, bias=True)
        conv(name + ".fc2", pl // 16, pl, 1, 1, bias=True)
    lin = nn.Linear(512, 10)
    params["linear.weight"] = lin.weight.detach().clone()
    params["linear.bias"] = lin.bias.detach().clone()
    return params, buffers


def senet_forward(x, params, buffers, training=True, num_blocks=RESNET18_BLOCKS, taps=None):
    out = F.conv2d(x, params["conv1.weight"], padding=1)
    if taps is not None:
        taps["conv1"] = out
    out = F.relu(_bn(out, "bn1", params, buffers, training))                               # senet.py:104
    for name, cin, pl, s in _senet_blocks(num_blocks):
        xin = out
        o = F.relu(_bn(xin, name + ".bn1", params, buffers, training))                     # :63
        sc = F.conv2d(o, params[name + ".shortcut.0.weight"], stride=s) if (name + ".shortcut.0.weight") in params \
            else xin                                                                       # :64
        o = F.conv2d(o, params[name + ".conv1.weight"], stride=s, padding=1)               # :65
        if taps is not None:
            taps[name + ".conv1"] = o
        o = F.conv2d(F.relu(_bn(o, name + ".bn2", params, buffers, training)), params[name + ".conv2.weight"],
                     padding=1)                                                            # :66
        if taps is not None:
            taps[name + ".conv2"] = o
        w = F.avg_pool2d(o, o.size(2))                                                     # :69
        w = F.relu(F.conv2d(w, params[name + ".fc1.weight"], params[name + ".fc1.bias"]))  # :70
        w = torch.sigmoid(F.conv2d(w, params[name + ".fc2.weight"], params[name + ".fc2.bias"]))   # :71
        out = o * w + sc                                                                   # :73-75
    out = F.avg_pool2d(out, 4).view(out.size(0), -1)                                       # :109-110
    return F.linear(out, params["linear.weight"], params["linear.bias"])


def build_regnety_400mf_params(seed):
    return build_regnetx_params(seed, "RegNetY_400MF")


def regnety_400mf_forward(x, params, buffers, training=True, taps=None):
    return regnetx_forward(x, params, buffers, training, "RegNetY_400MF", taps)


# ------------------------------------------------------------------------------------------------
# GoogLeNet (models/googlenet.py)
# ------------------------------------------------------------------------------------------------
GOOGLENET_CFG = [("a3", 192, 64, 96, 128, 16, 32, 32), ("b3", 256, 128, 128, 192, 32, 96, 64), "maxpool",
                 ("a4", 480, 192, 96, 208, 16, 48, 64), ("b4", 512, 160, 112, 224, 24, 64, 64),
                 ("c4", 512, 128, 128, 256, 24, 64, 64), ("d4", 512, 112, 144, 288, 32, 64, 64),
                 ("e4", 528, 256, 160, 320, 32, 128, 128), "maxpool",
                 ("a5", 832, 256, 160, 320, 32, 128, 128), ("b5", 832, 384, 192, 384, 48, 128, 128)]   # googlenet.py:64-78


def build_googlenet_params(seed):
    """Seeded parameters equal to `torch.manual_seed(seed); models.googlenet.GoogLeNet()`: pre_layers, then per
    Inception (googlenet.py:8-47) the branches b1, b2, b3, b4 in order, every conv WITH bias; Linear last."""
    torch.manual_seed(seed)
    params, buffers = OrderedDict(), OrderedDict()

    def conv_bn(cname, bname, cin, cout, k):
        m = nn.Conv2d(cin, cout, kernel_size=k, padding=k // 2)
        params[cname + ".weight"] = m.weight.detach().clone()
        params[cname + ".bias"] = m.bias.detach().clone()
        b = nn.BatchNorm2d(cout)
        params[bname + ".weight"] = b.weight.detach().clone()
        params[bname + ".bias"] = b.bias.detach().clone()
        buffers[bname + ".running_mean"] = b.running_mean.clone()
        buffers[bname + ".running_var"] = b.running_var.clone()
        buffers[bname + ".num_batches_tracked"] = b.num_batches_tracked.clone()

    conv_bn("pre_layers.0", "pre_layers.1", 3, 192, 3)
    for m in GOOGLENET_CFG:
        if m == "maxpool":
            continue
        name, cin, n1, n3r, n3, n5r, n5, npool = m
        conv_bn(name + ".b1.0", name + ".b1.1", cin, n1, 1)
        conv_bn(name + ".b2.0", name + ".b2.1", cin, n3r, 1)
        conv_bn(name + ".b2.3", name + ".b2.4", n3r, n3, 3)
        conv_bn(name + ".b3.0", name + ".b3.1", cin, n5r, 1)
        conv_bn(name + ".b3.3", name + ".b3.4", n5r, n5, 3)
        conv_bn(name + ".b3.6", name + ".b3.7", n5, n5, 3)
        conv_bn(name + ".b4.1", name + ".b4.2", cin, npool, 1)
    lin = nn.Linear(1024, 10)
    params["linear.weight"] = lin.weight.detach().clone()
    params["linear.bias"] = lin.bias.detach().clone()
    return params, buffers


def googlenet_forward(x, params, buffers, training=True, taps=None):
    def cbr(t, cname, bname, k):
        y = F.conv2d(t, params[cname + ".weight"], params[cname + ".bias"], padding=k // 2)
        if taps is not None:
            taps[cname] = y
        return F.relu(_bn(y, bname, params, buffers, training))
    out = cbr(x, "pre_layers.0", "pre_layers.1", 3)                                   # googlenet.py:82
    for m in GOOGLENET_CFG:
        if m == "maxpool":
            out = F.max_pool2d(out, 3, stride=2, padding=1)                           # :85, :91
            continue
        name = m[0]
        y1 = cbr(out, name + ".b1.0", name + ".b1.1", 1)                              # :53
        y2 = cbr(cbr(out, name + ".b2.0", name + ".b2.1", 1), name + ".b2.3", name + ".b2.4", 3)     # :54
        y3 = cbr(cbr(cbr(out, name + ".b3.0", name + ".b3.1", 1), name + ".b3.3", name + ".b3.4", 3),
                 name + ".b3.6", name + ".b3.7", 3)                                   # :55
        y4 = cbr(F.max_pool2d(out, 3, stride=1, padding=1), name + ".b4.1", name + ".b4.2", 1)        # :56
        out = torch.cat([y1, y2, y3, y4], 1)                                          # :57
    out = F.avg_pool2d(out, 8, stride=1).view(out.size(0), -1)                        # :94-95
    return F.linear(out, params["linear.weight"], params["linear.bias"])


# ------------------------------------------------------------------------------------------------
# PreActResNet-50/101/152 (models/preact_resnet.py:36-61, PreActBottleneck)
# ------------------------------------------------------------------------------------------------
def _preact_bottleneck_blocks(num_blocks):
    cin = 64
    for li, (nb, planes) in enumerate(zip(num_blocks, RESNET_PLANES)):
        for bi, s in enumerate([1 if li == 0 else 2] + [1] * (nb - 1)):
            yield f"layer{li + 1}.{bi}", cin, planes, 4 * planes, s
            cin = 4 * planes


def build_preact_bottleneck_params(seed, num_blocks=RESNET50_BLOCKS):
    torch.manual_seed(seed)
    params, buffers = OrderedDict(), OrderedDict()

    def conv(name, cin, cout, k, stride):
        m = nn.Conv2d(cin, cout, kernel_size=k, stride=stride, padding=k // 2, bias=False)
        params[name + ".weight"] = m.weight.detach().clone()

    def bn(name, c):
        m = nn.BatchNorm2d(c)
        params[name + ".weight"] = m.weight.detach().clone()
        params[name + ".bias"] = m.bias.detach().clone()
        buffers[name + ".running_mean"] = m.running_mean.clone()
        buffers[name + ".running_var"] = m.running_var.clone()
        buffers[name + ".num_batches_tracked"] = m.num_batches_tracked.clone()

    conv("conv1", 3, 64, 3, 1)
    cin_final = 64
    for name, cin, planes, cout, s in _preact_bottleneck_blocks(num_blocks):
        bn(name + ".bn1", cin)                                 # preact_resnet.py:42
        conv(name + ".conv1", cin, planes, 1, 1)               # :43
        bn(name + ".bn2", planes)
        conv(name + ".conv2", planes, planes, 3, s)            # :45
        bn(name + ".bn3", planes)
        conv(name + ".conv3", planes, cout, 1, 1)              # :47
        if s != 1 or cin != cout:                              # :49-52
            conv(name + ".shortcut.0", cin, cout, 1, s)
        cin_final = cout
    lin = nn.Linear(cin_final, 10)
    params["linear.weight"] = lin.weight.detach().clone()
    params["linear.bias"] = lin.bias.detach().clone()
    return params, buffers


def preact_bottleneck_forward(x, params, buffers, training=True, num_blocks=RESNET50_BLOCKS, taps=None):
    out = F.conv2d(x, params["conv1.weight"], padding=1)                                    # :85
    if taps is not None:
        taps["conv1"] = out
    for name, cin, planes, cout, s in _preact_bottleneck_blocks(num_blocks):
        xin = out
        o = F.relu(_bn(xin, name + ".bn1", params, buffers, training))                      # :55
        sc = F.conv2d(o, params[name + ".shortcut.0.weight"], stride=s) if (name + ".shortcut.0.weight") in params \
            else xin                                                                        # :56
        o = F.conv2d(o, params[name + ".conv1.weight"])                                     # :57
        if taps is not None:
            taps[name + ".conv1"] = o
        o = F.conv2d(F.relu(_bn(o, name + ".bn2", params, buffers, training)), params[name + ".conv2.weight"], stride=s,
                     padding=1)                                                             # :58
        o = F.conv2d(F.relu(_bn(o, name + ".bn3", params, buffers, training)), params[name + ".conv3.weight"])   # :59
        out = o + sc                                                                        # :60
    out = F.avg_pool2d(out, 4).view(out.size(0), -1)
    return F.linear(out, params["linear.weight"], params["linear.bias"])
