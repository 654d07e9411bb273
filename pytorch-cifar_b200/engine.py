"""Execution engine: turns a network description (the composition the reference's models/*.py express
with nn.Conv2d / nn.BatchNorm2d / F.relu / `+=`) into a static list of kernel launches over
pre-allocated NHWC bf16 buffers and flat fp32 parameter / gradient arenas.

Everything here is host-side plumbing: torch supplies device memory and streams; all arithmetic is
done by the sm_100a kernels behind the C-ABI (`ops`). Because buffer addresses and launch order are
static, a whole training step can be captured into one CUDA graph (see trainer.py).

Reference anchors: models/resnet.py:38-51 (BasicBlock.forward), :132-140 (ResNet.forward);
main.py:99-105 (step order).
"""
from collections import OrderedDict

import os

import torch

from . import ops

BN_EPS = 1e-5        # nn.BatchNorm2d default, models/resnet.py:25
BN_MOMENTUM = 0.1    # nn.BatchNorm2d default


def _align(n, a):
    return (n + a - 1) // a * a


# --------------------------------------------------------------------------------------------
# Network descriptions. A spec is a dict:
#   params : OrderedDict name -> shape, in the reference's net.parameters() order
#   buffers: OrderedDict name -> shape (BN running stats; num_batches_tracked handled separately)
#   layout : model-family specific topology used by the plan builders below
# --------------------------------------------------------------------------------------------
def resnet_spec(num_blocks=(2, 2, 2, 2), planes=(64, 128, 256, 512), num_classes=10, bottleneck=False):
    """ResNet-18/34 (BasicBlock) or, with bottleneck=True, ResNet-50/101/152 (Bottleneck, expansion 4) exactly as
    models/resnet.py:16-93,96-160 composes them."""
    if bottleneck:
        return _bottleneck_spec(num_blocks, planes, num_classes)
    params, buffers, blocks = OrderedDict(), OrderedDict(), []

    def conv(name, cout, cin, k):
        params[name + ".weight"] = (cout, cin, k, k)

    def bn(name, c):
        params[name + ".weight"] = (c,)
        params[name + ".bias"] = (c,)
        buffers[name + ".running_mean"] = (c,)
        buffers[name + ".running_var"] = (c,)

    conv("conv1", 64, 3, 3)
    bn("bn1", 64)
    in_planes = 64
    for li, (nb, pl) in enumerate(zip(num_blocks, planes)):
        strides = [1 if li == 0 else 2] + [1] * (nb - 1)
        for bi, s in enumerate(strides):
            name = f"layer{li + 1}.{bi}"
            conv(name + ".conv1", pl, in_planes, 3)
            bn(name + ".bn1", pl)
            conv(name + ".conv2", pl, pl, 3)
            bn(name + ".bn2", pl)
            shortcut = s != 1 or in_planes != pl
            if shortcut:
                conv(name + ".shortcut.0", pl, in_planes, 1)
                bn(name + ".shortcut.1", pl)
            blocks.append(dict(name=name, cin=in_planes, cout=pl, stride=s, shortcut=shortcut))
            in_planes = pl
    params["linear.weight"] = (num_classes, in_planes)
    params["linear.bias"] = (num_classes,)
    return dict(family="resnet", params=params, buffers=buffers, blocks=blocks, num_classes=num_classes,
                feat=in_planes)


VGG_CFG = {   # models/vgg.py:6-11
    "VGG11": [64, "M", 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
    "VGG13": [64, 64, "M", 128, 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
    "VGG16": [64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512, "M"],
    "VGG19": [64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M", 512, 512, 512, 512, "M"],
}


def vgg_spec(name="VGG16", num_classes=10):
    """VGG as models/vgg.py:14-38 composes it: [Conv3x3(bias) -> BN -> ReLU]* with MaxPool2d(2,2) between
    stages, a no-op AvgPool2d(1,1), Linear(512, 10). Sequential indices give the state_dict names."""
    params, buffers, layers = OrderedDict(), OrderedDict(), []
    i, cin = 0, 3
    for x in VGG_CFG[name]:
        if x == "M":
            layers.append(("pool",))
            i += 1
            continue
        params[f"features.{i}.weight"] = (x, cin, 3, 3)
        params[f"features.{i}.bias"] = (x,)          # cancelled by the BatchNorm that follows
        params[f"features.{i + 1}.weight"] = (x,)
        params[f"features.{i + 1}.bias"] = (x,)
        buffers[f"features.{i + 1}.running_mean"] = (x,)
        buffers[f"features.{i + 1}.running_var"] = (x,)
        layers.append(("conv", f"features.{i}", f"features.{i + 1}", cin, x))
        cin = x
        i += 3
    params["classifier.weight"] = (num_classes, 512)
    params["classifier.bias"] = (num_classes,)
    return dict(family="vgg", params=params, buffers=buffers, layers=layers, num_classes=num_classes, feat=512)


def preact_resnet_spec(num_blocks=(2, 2, 2, 2), planes=(64, 128, 256, 512), num_classes=10, bottleneck=False):
    """PreActResNet-18/34 as models/preact_resnet.py:12-34,64-94 composes it; bottleneck=True: PreActResNet-50/101/152
    (PreActBottleneck, :36-61: bn1 -> relu -> conv1x1 -> bn2 -> relu -> conv3x3(stride) -> bn3 -> relu -> conv1x1 (x4),
    += shortcut)."""
    params, buffers, blocks = OrderedDict(), OrderedDict(), []
    params["conv1.weight"] = (64, 3, 3, 3)
    in_planes = 64
    if bottleneck:
        def bn(name, c):
            params[name + ".weight"] = (c,)
            params[name + ".bias"] = (c,)
            buffers[name + ".running_mean"] = (c,)
            buffers[name + ".running_var"] = (c,)
        for li, (nb, pl) in enumerate(zip(num_blocks, planes)):
            for bi, s in enumerate([1 if li == 0 else 2] + [1] * (nb - 1)):
                name, cout = f"layer{li + 1}.{bi}", 4 * pl
                bn(name + ".bn1", in_planes)
                params[name + ".conv1.weight"] = (pl, in_planes, 1, 1)
                bn(name + ".bn2", pl)
                params[name + ".conv2.weight"] = (pl, pl, 3, 3)
                bn(name + ".bn3", pl)
                params[name + ".conv3.weight"] = (cout, pl, 1, 1)
                shortcut = s != 1 or in_planes != cout
                if shortcut:
                    params[name + ".shortcut.0.weight"] = (cout, in_planes, 1, 1)
                blocks.append(dict(name=name, cin=in_planes, planes=pl, cout=cout, stride=s, shortcut=shortcut))
                in_planes = cout
        params["linear.weight"] = (num_classes, in_planes)
        params["linear.bias"] = (num_classes,)
        return dict(family="preact", params=params, buffers=buffers, blocks=blocks, num_classes=num_classes,
                    feat=in_planes)
    for li, (nb, pl) in enumerate(zip(num_blocks, planes)):
        strides = [1 if li == 0 else 2] + [1] * (nb - 1)
        for bi, s in enumerate(strides):
            name = f"layer{li + 1}.{bi}"
            for bn, c in ((".bn1", in_planes),):
                params[name + bn + ".weight"] = (c,)
                params[name + bn + ".bias"] = (c,)
                buffers[name + bn + ".running_mean"] = (c,)
                buffers[name + bn + ".running_var"] = (c,)
            params[name + ".conv1.weight"] = (pl, in_planes, 3, 3)
            params[name + ".bn2.weight"] = (pl,)
            params[name + ".bn2.bias"] = (pl,)
            buffers[name + ".bn2.running_mean"] = (pl,)
            buffers[name + ".bn2.running_var"] = (pl,)
            params[name + ".conv2.weight"] = (pl, pl, 3, 3)
            shortcut = s != 1 or in_planes != pl
            if shortcut:
                params[name + ".shortcut.0.weight"] = (pl, in_planes, 1, 1)
            blocks.append(dict(name=name, cin=in_planes, cout=pl, stride=s, shortcut=shortcut))
            in_planes = pl
    params["linear.weight"] = (num_classes, in_planes)
    params["linear.bias"] = (num_classes,)
    return dict(family="preact", params=params, buffers=buffers, blocks=blocks, num_classes=num_classes,
                feat=in_planes)


MOBILENETV2_CFG = [(1, 16, 1, 1), (6, 24, 2, 1), (6, 32, 3, 2), (6, 64, 4, 2), (6, 96, 3, 1), (6, 160, 3, 2),
                   (6, 320, 1, 1)]      # (expansion, out_planes, num_blocks, stride), models/mobilenetv2.py:42-48


def mobilenetv2_spec(num_classes=10):
    """MobileNetV2 as models/mobilenetv2.py:11-77 composes it (CIFAR strides, F.relu not ReLU6)."""
    params, buffers, blocks, depthwise = OrderedDict(), OrderedDict(), [], set()

    def bn(name, c):
        params[name + ".weight"] = (c,)
        params[name + ".bias"] = (c,)
        buffers[name + ".running_mean"] = (c,)
        buffers[name + ".running_var"] = (c,)

    params["conv1.weight"] = (32, 3, 3, 3)
    bn("bn1", 32)
    cin, i = 32, 0
    for expansion, cout, nb, stride in MOBILENETV2_CFG:
        for s in [stride] + [1] * (nb - 1):
            name, planes = f"layers.{i}", expansion * cin
            params[name + ".conv1.weight"] = (planes, cin, 1, 1)
            bn(name + ".bn1", planes)
            params[name + ".conv2.weight"] = (planes, 1, 3, 3)
            depthwise.add(name + ".conv2.weight")
            bn(name + ".bn2", planes)
            params[name + ".conv3.weight"] = (cout, planes, 1, 1)
            bn(name + ".bn3", cout)
            shortcut = s == 1 and cin != cout
            if shortcut:
                params[name + ".shortcut.0.weight"] = (cout, cin, 1, 1)
                bn(name + ".shortcut.1", cout)
            blocks.append(dict(name=name, cin=cin, cout=cout, planes=planes, stride=s, shortcut=shortcut))
            cin = cout
            i += 1
    params["conv2.weight"] = (1280, 320, 1, 1)
    bn("bn2", 1280)
    params["linear.weight"] = (num_classes, 1280)
    params["linear.bias"] = (num_classes,)
    return dict(family="mobilenetv2", params=params, buffers=buffers, blocks=blocks, num_classes=num_classes,
                feat=1280, depthwise=depthwise, head_prefixes=("linear", "conv2", "bn2"))


def simple_dla_spec(num_classes=10):
    """SimpleDLA as models/dla_simple.py:59-116 composes it — the network main.py:71 constructs: three plain
    conv3x3 + BN + ReLU layers (3->16->16->32), then four aggregation trees of BasicBlocks (64 level 1; 128 and 256
    level 2, stride 2; 512 level 1, stride 2) and Linear(512, 10). A Tree registers its Root first, then the left
    and right sub-trees (dla_simple.py:61-70), which fixes the parameter order."""
    params, buffers = OrderedDict(), OrderedDict()

    def bn(name, c):
        params[name + ".weight"] = (c,)
        params[name + ".bias"] = (c,)
        buffers[name + ".running_mean"] = (c,)
        buffers[name + ".running_var"] = (c,)

    def block(name, cin, cout, stride):
        params[name + ".conv1.weight"] = (cout, cin, 3, 3)
        bn(name + ".bn1", cout)
        params[name + ".conv2.weight"] = (cout, cout, 3, 3)
        bn(name + ".bn2", cout)
        sc = stride != 1 or cin != cout
        if sc:
            params[name + ".shortcut.0.weight"] = (cout, cin, 1, 1)
            bn(name + ".shortcut.1", cout)
        return dict(kind="block", name=name, cin=cin, cout=cout, stride=stride, shortcut=sc)

    def tree(name, cin, cout, level, stride):
        params[name + ".root.conv.weight"] = (cout, 2 * cout, 1, 1)
        bn(name + ".root.bn", cout)
        if level == 1:
            left = block(name + ".left_tree", cin, cout, stride)
            right = block(name + ".right_tree", cout, cout, 1)
        else:
            left = tree(name + ".left_tree", cin, cout, level - 1, stride)
            right = tree(name + ".right_tree", cout, cout, level - 1, 1)
        return dict(kind="tree", name=name, cin=cin, cout=cout, stride=stride, left=left, right=right)

    plain = []
    for name, cin, cout in (("base", 3, 16), ("layer1", 16, 16), ("layer2", 16, 32)):
        params[name + ".0.weight"] = (cout, cin, 3, 3)
        bn(name + ".1", cout)
        plain.append((name, cin, cout))
    trees = [tree("layer3", 32, 64, 1, 1), tree("layer4", 64, 128, 2, 2), tree("layer5", 128, 256, 2, 2),
             tree("layer6", 256, 512, 1, 2)]
    params["linear.weight"] = (num_classes, 512)
    params["linear.bias"] = (num_classes,)
    catn = {n: 2 for n in params if n.endswith(".root.conv.weight")}
    return dict(family="dla", params=params, buffers=buffers, plain=plain, trees=trees, num_classes=num_classes,
                feat=512, catn=catn)


def senet_spec(num_blocks=(2, 2, 2, 2), planes=(64, 128, 256, 512), num_classes=10):
    """SENet18 as models/senet.py:44-117 composes it: conv3x3 stem + BN + ReLU, then PreActBlocks with
    squeeze-and-excitation — out = relu(bn1(x)); sc = shortcut(out) | x; out = conv2(relu(bn2(conv1(out))));
    w = sigmoid(fc2(relu(fc1(avg_pool(out))))); out = out * w + sc — avg_pool2d(4), Linear(512, 10).
    fc1 / fc2 are 1x1 nn.Conv2d WITH bias (senet.py:59-60): `se_fc` lists their weights, which are not conv operands."""
    params, buffers, blocks, se_fc = OrderedDict(), OrderedDict(), [], set()

    def bn(name, c):
        params[name + ".weight"] = (c,)
        params[name + ".bias"] = (c,)
        buffers[name + ".running_mean"] = (c,)
        buffers[name + ".running_var"] = (c,)

    params["conv1.weight"] = (64, 3, 3, 3)
    bn("bn1", 64)
    cin = 64
    for li, (nb, pl) in enumerate(zip(num_blocks, planes)):
        for bi, s in enumerate([1 if li == 0 else 2] + [1] * (nb - 1)):
            name = f"layer{li + 1}.{bi}"
            bn(name + ".bn1", cin)
            params[name + ".conv1.weight"] = (pl, cin, 3, 3)
            bn(name + ".bn2", pl)
            params[name + ".conv2.weight"] = (pl, pl, 3, 3)
            shortcut = s != 1 or cin != pl
            if shortcut:
                params[name + ".shortcut.0.weight"] = (pl, cin, 1, 1)
            params[name + ".fc1.weight"] = (pl // 16, pl, 1, 1)
            params[name + ".fc1.bias"] = (pl // 16,)
            params[name + ".fc2.weight"] = (pl, pl // 16, 1, 1)
            params[name + ".fc2.bias"] = (pl,)
            se_fc |= {name + ".fc1.weight", name + ".fc2.weight"}
            blocks.append(dict(name=name, cin=cin, cout=pl, stride=s, shortcut=shortcut))
            cin = pl
    params["linear.weight"] = (num_classes, cin)
    params["linear.bias"] = (num_classes,)
    return dict(family="senet", params=params, buffers=buffers, blocks=blocks, num_classes=num_classes, feat=cin,
                se_fc=se_fc)


GOOGLENET_CFG = [   # (name, in_planes, n1x1, n3x3red, n3x3, n5x5red, n5x5, pool_planes), models/googlenet.py:64-78
    ("a3", 192, 64, 96, 128, 16, 32, 32), ("b3", 256, 128, 128, 192, 32, 96, 64), "maxpool",
    ("a4", 480, 192, 96, 208, 16, 48, 64), ("b4", 512, 160, 112, 224, 24, 64, 64), ("c4", 512, 128, 128, 256, 24, 64, 64),
    ("d4", 512, 112, 144, 288, 32, 64, 64), ("e4", 528, 256, 160, 320, 32, 128, 128), "maxpool",
    ("a5", 832, 256, 160, 320, 32, 128, 128), ("b5", 832, 384, 192, 384, 48, 128, 128)]


def googlenet_spec(num_classes=10):
    """GoogLeNet as models/googlenet.py:7-99 composes it: conv3x3(3 -> 192, bias) + BN + ReLU, nine Inception modules —
    four branches (1x1; 1x1 -> 3x3; 1x1 -> 3x3 -> 3x3; MaxPool(3, 1, 1) -> 1x1), every conv with a bias and followed by
    BN + ReLU, outputs concatenated along the channels — with MaxPool(3, 2, 1) between the stages, AvgPool(8),
    Linear(1024, 10). The concatenation is never written: `cat_src` lists, for each 1x1 conv that reads one, the widths
    of the tensors it is made of (the four branch outputs of the module before)."""
    params, buffers, modules, cat_src = OrderedDict(), OrderedDict(), [], {}

    def conv_bn(cname, bname, cout, cin, k):
        params[cname + ".weight"] = (cout, cin, k, k)
        params[cname + ".bias"] = (cout,)
        params[bname + ".weight"] = (cout,)
        params[bname + ".bias"] = (cout,)
        buffers[bname + ".running_mean"] = (cout,)
        buffers[bname + ".running_var"] = (cout,)

    conv_bn("pre_layers.0", "pre_layers.1", 192, 3, 3)
    src = [192]
    for m in GOOGLENET_CFG:
        if m == "maxpool":
            modules.append("maxpool")
            continue
        name, cin, n1, n3r, n3, n5r, n5, npool = m
        assert cin == sum(src), name
        conv_bn(name + ".b1.0", name + ".b1.1", n1, cin, 1)
        conv_bn(name + ".b2.0", name + ".b2.1", n3r, cin, 1)
        conv_bn(name + ".b2.3", name + ".b2.4", n3, n3r, 3)
        conv_bn(name + ".b3.0", name + ".b3.1", n5r, cin, 1)
        conv_bn(name + ".b3.3", name + ".b3.4", n5, n5r, 3)
        conv_bn(name + ".b3.6", name + ".b3.7", n5, n5, 3)
        conv_bn(name + ".b4.1", name + ".b4.2", npool, cin, 1)
        for c in (".b1.0", ".b2.0", ".b3.0", ".b4.1"):
            cat_src[name + c + ".weight"] = list(src)
        modules.append(dict(name=name, n1=n1, n3r=n3r, n3=n3, n5r=n5r, n5=n5, npool=npool))
        src = [n1, n3, n5, npool]
    params["linear.weight"] = (num_classes, sum(src))
    params["linear.bias"] = (num_classes,)
    return dict(family="googlenet", params=params, buffers=buffers, modules=modules, num_classes=num_classes,
                feat=sum(src), cat_src=cat_src)


def dla_spec(num_classes=10):
    """DLA as models/dla.py:53-125 composes it: the three plain conv layers of SimpleDLA, then trees whose level-2
    form (layer4, layer5) is  xs = [prev_root(x)]; x = level_1(x); x = left_node(x); x = right_node(x);
    root(xs + the three) — a Root over FOUR concatenated tensors — where level_1 is a level-1 tree (root over
    [left_node, right_node]). Registration order (dla.py:57-69): root, level_1, prev_root, left_node, right_node."""
    params, buffers, catn = OrderedDict(), OrderedDict(), {}

    def bn(name, c):
        params[name + ".weight"] = (c,)
        params[name + ".bias"] = (c,)
        buffers[name + ".running_mean"] = (c,)
        buffers[name + ".running_var"] = (c,)

    def block(name, cin, cout, stride):
        params[name + ".conv1.weight"] = (cout, cin, 3, 3)
        bn(name + ".bn1", cout)
        params[name + ".conv2.weight"] = (cout, cout, 3, 3)
        bn(name + ".bn2", cout)
        sc = stride != 1 or cin != cout
        if sc:
            params[name + ".shortcut.0.weight"] = (cout, cin, 1, 1)
            bn(name + ".shortcut.1", cout)
        return dict(kind="block", name=name, cin=cin, cout=cout, stride=stride, shortcut=sc)

    def root(name, n_src, cout):
        params[name + ".root.conv.weight"] = (cout, n_src * cout, 1, 1)
        catn[name + ".root.conv.weight"] = n_src
        bn(name + ".root.bn", cout)

    def tree(name, cin, cout, level, stride):
        if level == 1:
            root(name, 2, cout)
            left = block(name + ".left_node", cin, cout, stride)
            right = block(name + ".right_node", cout, cout, 1)
            return dict(kind="tree", name=name, cin=cin, cout=cout, stride=stride, left=left, right=right)
        assert level == 2
        root(name, 4, cout)
        level_1 = tree(name + ".level_1", cin, cout, 1, stride)
        prev_root = block(name + ".prev_root", cin, cout, stride)
        left = block(name + ".left_node", cout, cout, 1)
        right = block(name + ".right_node", cout, cout, 1)
        return dict(kind="tree2", name=name, cin=cin, cout=cout, stride=stride, level_1=level_1, prev_root=prev_root,
                    left=left, right=right)

    plain = []
    for name, cin, cout in (("base", 3, 16), ("layer1", 16, 16), ("layer2", 16, 32)):
        params[name + ".0.weight"] = (cout, cin, 3, 3)
        bn(name + ".1", cout)
        plain.append((name, cin, cout))
    trees = [tree("layer3", 32, 64, 1, 1), tree("layer4", 64, 128, 2, 2), tree("layer5", 128, 256, 2, 2),
             tree("layer6", 256, 512, 1, 2)]
    params["linear.weight"] = (num_classes, 512)
    params["linear.bias"] = (num_classes,)
    return dict(family="dla", params=params, buffers=buffers, plain=plain, trees=trees, num_classes=num_classes,
                feat=512, catn=catn)


MOBILENET_CFG = [64, (128, 2), 128, (256, 2), 256, (512, 2), 512, 512, 512, 512, 512, (1024, 2), 1024]   # models/mobilenet.py:27


def mobilenet_spec(num_classes=10):
    """MobileNet (v1) as models/mobilenet.py:11-52 composes it: conv3x3(3->32) + BN + ReLU, 13 blocks of depthwise 3x3
    (stride 1 or 2) + BN + ReLU -> pointwise 1x1 + BN + ReLU, avg_pool2d(2), Linear(1024, 10)."""
    params, buffers, blocks, depthwise = OrderedDict(), OrderedDict(), [], set()

    def bn(name, c):
        params[name + ".weight"] = (c,)
        params[name + ".bias"] = (c,)
        buffers[name + ".running_mean"] = (c,)
        buffers[name + ".running_var"] = (c,)

    params["conv1.weight"] = (32, 3, 3, 3)
    bn("bn1", 32)
    cin = 32
    for i, x in enumerate(MOBILENET_CFG):
        cout, s = (x, 1) if isinstance(x, int) else x
        name = f"layers.{i}"
        params[name + ".conv1.weight"] = (cin, 1, 3, 3)
        depthwise.add(name + ".conv1.weight")
        bn(name + ".bn1", cin)
        params[name + ".conv2.weight"] = (cout, cin, 1, 1)
        bn(name + ".bn2", cout)
        blocks.append(dict(name=name, cin=cin, cout=cout, stride=s))
        cin = cout
    params["linear.weight"] = (num_classes, cin)
    params["linear.bias"] = (num_classes,)
    return dict(family="mobilenet", params=params, buffers=buffers, blocks=blocks, num_classes=num_classes, feat=cin,
                depthwise=depthwise)


def _pad64(c):
    return _align(c, 64)


def param_offsets(spec):
    """Element offsets of every parameter in the flat fp32 arenas (16-byte aligned), reference order.
    Per-channel vectors (BatchNorm weight / bias) are ALLOCATED padded to the 64-channel layout width the kernels
    use; the padding stays zero (zero gradient, zero weight decay), the named views cover the true shape."""
    p_off, off = OrderedDict(), 0
    for name, shape in spec["params"].items():
        n = 1
        for s in shape:
            n *= s
        p_off[name] = (off, n, shape)
        alloc = _pad64(n) if len(shape) == 1 and not name.startswith(("linear", "classifier")) else n
        off = _align(off + alloc, 4)
    return p_off, _align(off, 4)


def backward_groups(spec):
    """Parameter-name groups in the order the backward plan completes them (head, blocks reversed, stem)."""
    first_conv = next(n for n, sh in spec["params"].items() if len(sh) == 4)[:-len(".weight")]

    def key(name):
        if spec["family"] in ("dla", "googlenet"):   # one backward op per top-level layer (a tree / an Inception module)
            top = name.split(".")[0]
            return "head" if top == "linear" else ("stem" if top in ("base", "pre_layers") else top)
        if name.startswith("layer"):
            return ".".join(name.split(".")[:2])
        if name.startswith("features."):
            idx = int(name.split(".")[1])
            owner = idx if f"features.{idx}.weight" in spec["params"] and len(spec["params"][f"features.{idx}.weight"]) == 4 \
                else idx - 1
            return "stem" if f"features.{owner}" == first_conv else f"features.{owner}"
        return "head" if name.startswith(spec.get("head_prefixes", ("linear", "classifier"))) else "stem"
    groups = OrderedDict()
    for name in spec["params"]:
        groups.setdefault(key(name), []).append(name)
    order = ["head"] + [k for k in reversed(list(groups)) if k not in ("head", "stem")] + ["stem"]
    return [(k, groups[k]) for k in order]


def plan_grad_buckets(spec, bucket_bytes):
    """Cut the gradient arena into contiguous buckets in backward order, each closed once it holds at least
    bucket_bytes (the DDP Reducer's rule, reducer.hpp:30-31, over a pre-flattened arena).
    Returns (buckets, ready_after): buckets = [(start, end)] element ranges; ready_after maps the index of a
    backward op (0 = head, 1.. = blocks in reverse, last = stem) to the bucket that is complete after it."""
    p_off, _ = param_offsets(spec)
    buckets, ready_after = [], {}
    start = end = None
    for pos, (key, names) in enumerate(backward_groups(spec)):
        a = min(p_off[n][0] for n in names)
        b = _align(max(p_off[n][0] + p_off[n][1] for n in names), 4)   # include the alignment padding
        if end is None:
            end = b
        start = a
        if (end - start) * 4 >= bucket_bytes:
            ready_after[pos] = len(buckets)
            buckets.append((start, end))
            end = None
    if end is not None:
        ready_after[len(backward_groups(spec)) - 1] = len(buckets)
        buckets.append((start, end))
    return buckets, ready_after


def init_params(spec, seed=0):
    """Seeded initial parameters and BatchNorm buffers of `spec`, equal bit for bit to
    `torch.manual_seed(seed); net = <reference constructor>()` (e.g. models/resnet.py:143-144): the reference's
    constructors create their layers in `net.parameters()` order and each layer draws from torch's global CPU
    generator in nn.Conv2d / nn.Linear.reset_parameters (kaiming_uniform_(a=sqrt(5)) for the weight, then
    U(-1/sqrt(fan_in), 1/sqrt(fan_in)) for a bias); BatchNorm draws nothing (weight 1, bias 0, running_var 1).
    Returns (params, buffers): OrderedDicts of CPU fp32 tensors keyed and shaped like the reference's state_dict
    (without num_batches_tracked)."""
    import math
    torch.manual_seed(seed)
    params = OrderedDict()
    shapes = spec["params"]
    for name, shape in shapes.items():
        if name.endswith(".weight") and len(shape) >= 2:
            w = torch.empty(*shape)
            torch.nn.init.kaiming_uniform_(w, a=math.sqrt(5))
            params[name] = w
        elif name.endswith(".weight"):
            params[name] = torch.ones(*shape)
        else:                                              # a bias: conv / linear (drawn) or BatchNorm (zero)
            wshape = shapes[name[:-len("bias")] + "weight"]
            if len(wshape) >= 2:
                fan_in = 1
                for d in wshape[1:]:
                    fan_in *= d
                bound = 1.0 / math.sqrt(fan_in)
                params[name] = torch.empty(*shape).uniform_(-bound, bound)
            else:
                params[name] = torch.zeros(*shape)
    buffers = OrderedDict((name, torch.ones(*shape) if name.endswith("running_var") else torch.zeros(*shape))
                          for name, shape in spec["buffers"].items())
    return params, buffers


def _bn_views(u, eng, bn_name, C):
    """Padded-width views of a BatchNorm's parameters / statistics for the kernels."""
    u.gamma, u.beta = eng.param_vec(bn_name + ".weight", C), eng.param_vec(bn_name + ".bias", C)
    u.dgamma, u.dbeta = eng.grad_vec(bn_name + ".weight", C), eng.grad_vec(bn_name + ".bias", C)
    u.rm, u.rv = eng.buffer_vec(bn_name + ".running_mean", C), eng.buffer_vec(bn_name + ".running_var", C)
    u.scale, u.shift, u.mean, u.rstd = (eng.new_f32(C) for _ in range(4))
    if not hasattr(u, "fsum"):
        u.fsum = eng.new_stat(2 * C, forward=True)     # [sum | sum of squares] of y, accumulated by its producer
    u.bsum = eng.new_stat(2 * C, forward=False)        # [sum dz | sum dz*y] of the backward pass
    u.make_stat_rec()


def resnet50_spec(num_classes=10):
    return resnet_spec((3, 4, 6, 3), num_classes=num_classes, bottleneck=True)


def _bottleneck_spec(num_blocks, planes, num_classes):
    params, buffers, blocks = OrderedDict(), OrderedDict(), []

    def bn(name, c):
        params[name + ".weight"] = (c,)
        params[name + ".bias"] = (c,)
        buffers[name + ".running_mean"] = (c,)
        buffers[name + ".running_var"] = (c,)

    params["conv1.weight"] = (64, 3, 3, 3)
    bn("bn1", 64)
    cin = 64
    for li, (nb, pl) in enumerate(zip(num_blocks, planes)):
        for bi, s in enumerate([1 if li == 0 else 2] + [1] * (nb - 1)):
            name, cout = f"layer{li + 1}.{bi}", 4 * pl
            params[name + ".conv1.weight"] = (pl, cin, 1, 1)
            bn(name + ".bn1", pl)
            params[name + ".conv2.weight"] = (pl, pl, 3, 3)
            bn(name + ".bn2", pl)
            params[name + ".conv3.weight"] = (cout, pl, 1, 1)
            bn(name + ".bn3", cout)
            shortcut = s != 1 or cin != cout
            if shortcut:
                params[name + ".shortcut.0.weight"] = (cout, cin, 1, 1)
                bn(name + ".shortcut.1", cout)
            blocks.append(dict(name=name, cin=cin, planes=pl, cout=cout, stride=s, shortcut=shortcut))
            cin = cout
    params["linear.weight"] = (num_classes, cin)
    params["linear.bias"] = (num_classes,)
    return dict(family="bottleneck", params=params, buffers=buffers, blocks=blocks, num_classes=num_classes, feat=cin)


def resnext_spec(num_blocks=(3, 3, 3), cardinality=2, bottleneck_width=64, num_classes=10):
    """ResNeXt-29 as models/resnext.py:10-71 composes it: a 1x1 stem conv (3 -> 64) + BN + ReLU, three stages of
    `Block`s — conv1x1 -> BN -> ReLU -> grouped conv3x3 (groups = cardinality, stride) -> BN -> ReLU -> conv1x1 (x2)
    -> BN -> += shortcut (1x1 conv + BN where the shape changes) -> ReLU — with the group width doubling per stage,
    avg_pool2d(8), Linear(cardinality * bottleneck_width * 8, 10). `groups` maps the grouped convs' weights to g."""
    params, buffers, blocks, groups = OrderedDict(), OrderedDict(), [], {}

    def bn(name, c):
        params[name + ".weight"] = (c,)
        params[name + ".bias"] = (c,)
        buffers[name + ".running_mean"] = (c,)
        buffers[name + ".running_var"] = (c,)

    params["conv1.weight"] = (64, 3, 1, 1)
    bn("bn1", 64)
    cin, bw = 64, bottleneck_width
    for li, nb in enumerate(num_blocks):
        for bi, s in enumerate([1 if li == 0 else 2] + [1] * (nb - 1)):
            name, gw = f"layer{li + 1}.{bi}", cardinality * bw
            cout = 2 * gw
            params[name + ".conv1.weight"] = (gw, cin, 1, 1)
            bn(name + ".bn1", gw)
            params[name + ".conv2.weight"] = (gw, gw // cardinality, 3, 3)
            groups[name + ".conv2.weight"] = cardinality
            bn(name + ".bn2", gw)
            params[name + ".conv3.weight"] = (cout, gw, 1, 1)
            bn(name + ".bn3", cout)
            shortcut = s != 1 or cin != cout
            if shortcut:
                params[name + ".shortcut.0.weight"] = (cout, cin, 1, 1)
                bn(name + ".shortcut.1", cout)
            blocks.append(dict(name=name, cin=cin, planes=gw, cout=cout, stride=s, shortcut=shortcut))
            cin = cout
        bw *= 2
    params["linear.weight"] = (num_classes, cin)
    params["linear.bias"] = (num_classes,)
    return dict(family="bottleneck", params=params, buffers=buffers, blocks=blocks, num_classes=num_classes, feat=cin,
                groups=groups, stem_ksize=1)


REGNET_CFG = {   # models/regnet.py:113-146
    "RegNetX_200MF": dict(depths=(1, 1, 4, 7), widths=(24, 56, 152, 368), strides=(1, 1, 2, 2), group_width=8, se_ratio=0),
    "RegNetX_400MF": dict(depths=(1, 2, 7, 12), widths=(32, 64, 160, 384), strides=(1, 1, 2, 2), group_width=16, se_ratio=0),
    "RegNetY_400MF": dict(depths=(1, 2, 7, 12), widths=(32, 64, 160, 384), strides=(1, 1, 2, 2), group_width=16,
                          se_ratio=0.25),
}


def regnetx_spec(name="RegNetX_200MF", num_classes=10):
    """RegNetX / RegNetY as models/regnet.py:29-110 composes them (bottleneck_ratio 1): conv3x3 stem (3 -> 64) + BN +
    ReLU, four stages of `Block`s — conv1x1 -> BN -> ReLU -> grouped conv3x3 (groups = width / group_width, stride) ->
    BN -> ReLU -> [RegNetY: squeeze-and-excitation with round(w_in / 4) hidden units, regnet.py:42-45] -> conv1x1 -> BN
    -> += shortcut (1x1 conv + BN where the shape changes) -> ReLU —, adaptive_avg_pool2d(1) over the final 8x8 map,
    Linear(widths[-1], 10)."""
    cfg = REGNET_CFG[name]
    params, buffers, blocks, groups, se_fc = OrderedDict(), OrderedDict(), [], {}, set()

    def bn(bname, c):
        params[bname + ".weight"] = (c,)
        params[bname + ".bias"] = (c,)
        buffers[bname + ".running_mean"] = (c,)
        buffers[bname + ".running_var"] = (c,)

    params["conv1.weight"] = (64, 3, 3, 3)
    bn("bn1", 64)
    cin = 64
    for li, (depth, width, stride) in enumerate(zip(cfg["depths"], cfg["widths"], cfg["strides"])):
        for bi in range(depth):
            bname, s, g = f"layer{li + 1}.{bi}", (stride if bi == 0 else 1), width // cfg["group_width"]
            params[bname + ".conv1.weight"] = (width, cin, 1, 1)
            bn(bname + ".bn1", width)
            params[bname + ".conv2.weight"] = (width, width // g, 3, 3)
            if g > 1:
                groups[bname + ".conv2.weight"] = g
            bn(bname + ".bn2", width)
            se = 0
            if cfg["se_ratio"] > 0:
                se = int(round(cin * cfg["se_ratio"]))                   # regnet.py:44: w_in, not w_b
                params[bname + ".se.se1.weight"] = (se, width, 1, 1)
                params[bname + ".se.se1.bias"] = (se,)
                params[bname + ".se.se2.weight"] = (width, se, 1, 1)
                params[bname + ".se.se2.bias"] = (width,)
                se_fc |= {bname + ".se.se1.weight", bname + ".se.se2.weight"}
            params[bname + ".conv3.weight"] = (width, width, 1, 1)
            bn(bname + ".bn3", width)
            shortcut = s != 1 or cin != width
            if shortcut:
                params[bname + ".shortcut.0.weight"] = (width, cin, 1, 1)
                bn(bname + ".shortcut.1", width)
            blocks.append(dict(name=bname, cin=cin, planes=width, cout=width, stride=s, shortcut=shortcut, se=se))
            cin = width
    params["linear.weight"] = (num_classes, cin)
    params["linear.bias"] = (num_classes,)
    return dict(family="bottleneck", params=params, buffers=buffers, blocks=blocks, num_classes=num_classes, feat=cin,
                groups=groups, se_fc=se_fc)


def regnetx_400mf_spec(num_classes=10):
    return regnetx_spec("RegNetX_400MF", num_classes)


def regnety_400mf_spec(num_classes=10):
    return regnetx_spec("RegNetY_400MF", num_classes)


def preact_resnet50_spec(num_classes=10):
    return preact_resnet_spec((3, 4, 6, 3), num_classes=num_classes, bottleneck=True)


def resnext29_4x64d_spec(num_classes=10):
    return resnext_spec(cardinality=4, bottleneck_width=64, num_classes=num_classes)


def resnext29_32x4d_spec(num_classes=10):
    return resnext_spec(cardinality=32, bottleneck_width=4, num_classes=num_classes)


# measurement switch (A/B runs on one box): 0 = separate dgrad and bn_bwd_reduce launches
FUSE_DGRAD_BN = os.environ.get("B2C_FUSE_DGRAD_BN", "1") != "0"
# 0 = residual units take their ReLU mask from the activation instead of the forward pass's bit mask
RELU_BITS = os.environ.get("B2C_RELU_BITS", "1") != "0"
# 0 = conv weights / gradients stored OIHW in the arenas (scattered 4-byte wgrad atomics)
TCO_LAYOUT = os.environ.get("B2C_TCO_LAYOUT", "1") != "0"
# 1 = weight-gradient kernels run on a side stream, concurrently with the data-gradient chain that does not
# depend on them (fork after the BN backward that produced dy, join before the optimizer / the bucket's all-reduce)
ASYNC_WGRAD = os.environ.get("B2C_ASYNC_WGRAD", "1") != "0"
# 1 = a stride-1 depthwise conv reads the RAW output of the 1x1 conv before it and applies that conv's BatchNorm + ReLU
# while loading (b2c_dwconv_fprop_bnrelu / b2c_dwconv_wgrad_bnrelu): the expanded activation of a MobileNetV2 block —
# its widest tensor — is never written or re-read; 0 = separate b2c_bn_act_apply pass. Measured (batch 512,
# profiles/r2o_dw_fused_input_ab.txt): 8.25 ms per step unfused, 10.10 ms fused — the strip kernels are bound by their
# instruction issue, not by HBM, and the transform (fma + max + border select per loaded value, three loads per
# output) doubles their instruction count, which costs more than the saved 4 B per element. Off by default.
DW_FUSED_INPUT = os.environ.get("B2C_DW_FUSED_INPUT", "0") != "0"
# 1 = a grouped conv whose group width is a multiple of 64 (ResNeXt29_2x64d / 4x64d / 8x64d) runs one dense conv per
# group on channel slices; 0 = always the dense conv over block-diagonal operands (g x the FLOPs)
GROUP_SLICES = os.environ.get("B2C_GROUP_SLICES", "1") != "0"
# CTA-pair (tcgen05.mma.cta_group::2) fprop / dgrad kernels: "halo" = where the halo kernel runs (3x3 stride 1 on
# 32/16/8-pixel images), "all" = the per-tap kernel's shapes too, "0" = single-CTA kernels everywhere
CTA_PAIR = os.environ.get("B2C_CTA_PAIR", "all")


# 1 = BatchNorm statistics are accumulated into one row per BatchNorm (fp32 atomics in the producing kernel's epilogue)
# and finalized in the prologue of the pass that applies them (b2c_bn_act_apply_stats / b2c_bn_bwd_apply_stats): no
# bn_finalize / bn_bwd_finalize launches (40 of a ResNet-18 step's 156); 0 = per-CTA partial rows + finalize kernels
FUSED_FINALIZE = os.environ.get("B2C_FUSED_FINALIZE", "0") != "0"


class _BNOps:
    """BatchNorm forward apply / backward of a unit that has y, rows, cout and the _bn_views fields."""

    def make_stat_rec(self):
        C = self.cout
        self.stat_rec = ops.bn_stats(self.fsum[:C], self.fsum[C:], self.gamma, self.beta, self.rows, BN_EPS, BN_MOMENTUM,
                                     self.rm, self.rv, self.scale, self.shift, self.mean, self.rstd,
                                     conv_bias=getattr(self, "conv_bias", None))

    def bn_stats(self, eng, training):
        if training:
            ops.bn_finalize(self.ps, self.pq, self.stat_rows, self.cout, self.rows, self.gamma, self.beta,
                            BN_EPS, BN_MOMENTUM, self.rm, self.rv, self.scale, self.shift, self.mean, self.rstd,
                            conv_bias=getattr(self, "conv_bias", None))
        else:
            ops.bn_eval_scale_shift(self.cout, self.gamma, self.beta, self.rm, self.rv, BN_EPS, self.scale,
                                    self.shift, conv_bias=getattr(self, "conv_bias", None))

    def bn_scale_shift(self, eng, training):
        """Only the BatchNorm's scale / shift / saved statistics / running statistics: for a unit whose apply pass is
        folded into its consumer (DwUnit with src_unit)."""
        if FUSED_FINALIZE and training:      # statistics were accumulated into one row by the producer
            C = self.cout
            ops.bn_finalize(self.fsum[:C], self.fsum[C:], 1, C, self.rows, self.gamma, self.beta, BN_EPS, BN_MOMENTUM,
                            self.rm, self.rv, self.scale, self.shift, self.mean, self.rstd,
                            conv_bias=getattr(self, "conv_bias", None))
        else:
            self.bn_stats(eng, training)

    def bn_apply(self, eng, training, out, relu=True, res=None, res_unit=None, relu_bits=None):
        """out = act(bn(y) + residual), residual = `res` (identity shortcut) or res_unit's own BatchNorm of
        res_unit.y (conv shortcut) — models/resnet.py:47-50. With FUSED_FINALIZE (training) one launch finalizes
        the statistics of both BatchNorms and applies them; otherwise bn_stats() (for res_unit: issued by the
        caller, on the side stream) followed by the apply pass."""
        if FUSED_FINALIZE and training:
            ops.bn_act_apply_stats(self.y, self.stat_rec, out, self.rows, self.cout, relu=relu,
                                   res=res_unit.y if res_unit is not None else res,
                                   res_bn=res_unit.stat_rec if res_unit is not None else None, relu_bits=relu_bits)
            return
        self.bn_stats(eng, training)
        if res_unit is not None:
            ops.bn_act_apply(self.y, self.scale, self.shift, out, self.rows, self.cout, relu=relu, relu_bits=relu_bits,
                             res=res_unit.y, res_scale=res_unit.scale, res_shift=res_unit.shift)
        else:
            ops.bn_act_apply(self.y, self.scale, self.shift, out, self.rows, self.cout, relu=relu, relu_bits=relu_bits,
                             res=res)

    def bn_backward(self, eng, dout, act, dz_out, dy, add=None, relu_from_y=False, relu_bits=None):
        """dout: grad w.r.t. act(bn(y)) (or w.r.t. bn(y) when act is None). Writes dz (optional) and
        dy = dBN/dy (+ add). relu_from_y: the unit is a plain relu(bn(y)) and nothing else needs dz: the mask is
        recomputed from y with this BN's own scale/shift, so neither `act` is read nor dz written."""
        reduce = ops.bn_bwd_reduce_acc if FUSED_FINALIZE else ops.bn_bwd_reduce
        part = self.bsum if FUSED_FINALIZE else eng.bwd_partial
        if relu_from_y:
            reduce(dout, None, self.y, None, self.rows, self.cout, part, self.scale, self.shift)
        elif relu_bits is not None:        # mask saved by the forward apply pass: `act` is not read
            reduce(dout, None, self.y, dz_out, self.rows, self.cout, part, relu_bits=relu_bits)
        else:
            reduce(dout, act, self.y, dz_out, self.rows, self.cout, part)
        src = dz_out if (dz_out is not None and not relu_from_y) else dout
        self._bn_bwd_finish(eng, src, dy, ops.bn_bwd_partial_rows(self.rows, self.cout), add,
                            self.scale if relu_from_y else None, self.shift if relu_from_y else None)

    def bn_backward_from_partial(self, eng, dz, dy, prow, add=None):
        """Second half of bn_backward when the producer of dz already masked it and reduced it (dgrad_bn)."""
        self._bn_bwd_finish(eng, dz, dy, prow, add, None, None)

    def _bn_bwd_finish(self, eng, src, dy, prow, add, relu_scale, relu_shift):
        if FUSED_FINALIZE:
            ops.bn_bwd_apply_stats(src, self.y, self.bsum, self.rows, self.gamma, self.mean, self.rstd, self.dgamma,
                                   self.dbeta, dy, self.rows, self.cout, add=add, relu_scale=relu_scale,
                                   relu_shift=relu_shift)
            return
        ops.bn_bwd_finalize(eng.bwd_partial, prow, self.cout, self.rows, self.gamma, self.mean, self.rstd,
                            self.dgamma, self.dbeta, eng.coef_a, eng.coef_b, eng.coef_c)
        ops.bn_bwd_apply(src, self.y, eng.coef_a, eng.coef_b, eng.coef_c, dy, self.rows, self.cout, add=add,
                         relu_scale=relu_scale, relu_shift=relu_shift)


class ConvUnit(_BNOps):
    """One conv (+ the BatchNorm that follows it): descriptors, operand views, statistic buffers.
    Channel counts are the reference's; tensors are laid out with channels padded to multiples of 64
    (self.cin / self.cout are those layout widths, *_true the reference's)."""

    def __init__(self, eng, name, bn_name, x, cin, cout, k, stride, H, W, stem=False, private_stats=False):
        """stem: x is the stem operand built by the input kernels (im2col of the 3-channel image for k = 3, the image
        itself padded to 64 channels for k = 1). A grouped conv (spec["groups"]) runs as a dense cin -> cout conv over
        block-diagonal operands; its weight gradient goes through a dense scratch matrix (see wgrad)."""
        N = eng.N
        self.name, self.bn_name = name, bn_name
        self.groups = eng.spec.get("groups", {}).get(name + ".weight", 1)
        self.cin_true, self.cout_true = cin, cout
        cin_p, cout_p = (64 if stem else _pad64(cin)), _pad64(cout)
        self.cin, self.cout, self.k, self.stride, self.H, self.W = cin_p, cout_p, k, stride, H, W
        self.Ho, self.Wo = H // stride, W // stride
        self.rows = N * self.Ho * self.Wo
        self.stem = stem
        self.x = x
        self.y = eng.new_act(N, self.Ho, self.Wo, cout_p)
        if stem:
            # the 3-channel stem runs as a 1x1 GEMM over an im2col operand (k = 27 padded to 64)
            self.desc = ops.conv_desc(N, H, W, 64, cout_p, 1, 1, flags=512 if FUSED_FINALIZE else 0)
            if k == 3:
                self.desc_w = ops.conv_desc(N, H, W, 64, cout_p, 1, 1, cin_valid=27, kflat_cin=3, cout_valid=cout)
            else:
                self.desc_w = ops.conv_desc(N, H, W, 64, cout_p, 1, 1, cin_valid=3, cout_valid=cout)
        else:
            halo_shape = k == 3 and stride == 1 and H == W and W in (32, 16, 8)
            pair = CTA_PAIR == "all" or (CTA_PAIR == "halo" and halo_shape)
            self.desc = ops.conv_desc(N, H, W, cin_p, cout_p, k, stride,
                                      flags=(128 if pair else 0) | (512 if FUSED_FINALIZE else 0))
            tco_out = (name + ".weight") in eng.tco or self.groups > 1       # gradient written [tap][cin][cout]
            self.desc_w = ops.conv_desc(N, H, W, cin_p, cout_p, k, stride, cin_valid=cin, cout_valid=cout,
                                        flags=(32 | (128 if CTA_PAIR != "0" else 0)) if tco_out else 0)
        self.eng_scratch = eng.wgrad_scratch
        self.eng = eng
        self.sliced = self.groups > 1 and name in eng.group_ops
        if self.sliced:
            # one dense conv per group on channel slices of x / y (see _layout_params); statistics are accumulated
            # into this unit's own row (every group adds its slice), which forward() zero-fills
            g, cg_in, cg_out = self.groups, cin_p // self.groups, cout_p // self.groups
            eng.zero_stats_each_forward = True
            self.g_desc = ops.conv_desc(N, H, W, cg_in, cg_out, k, stride, flags=(128 if pair else 0) | 512,
                                        x_ld=cin_p, y_ld=cout_p)
            self.g_desc_w = ops.conv_desc(N, H, W, cg_in, cg_out, k, stride, x_ld=cin_p, y_ld=cout_p)
            self.g_wf = [eng.WF[a:a + sf[0] * sf[1] * sf[2]].view(sf) for a, _, sf, _ in eng.group_ops[name]]
            self.g_wd = [eng.WD[b:b + sd[0] * sd[1] * sd[2]].view(sd) for _, b, _, sd in eng.group_ops[name]]
            dwf = eng.grad_flat(name + ".weight")                      # [cout][cin/g][kh][kw]: group j = rows j*cout/g..
            per = cg_out * cg_in * k * k
            self.g_dw = [dwf[j * per:(j + 1) * per] for j in range(g)]
            self.g_in = [j * cg_in for j in range(g)]
            self.g_out = [j * cg_out for j in range(g)]
        elif self.groups > 1:
            eng.grouped_dense(k * k * cin_p * cout_p)
        # BatchNorm partial-sum scratch: shared by the units of the main chain (used strictly one after the other);
        # a shortcut conv that runs concurrently on the side stream gets its own
        if private_stats:
            self.ps = torch.zeros(ops.MAX_STAT_ROWS, _pad64(cout), device=eng.device)
            self.pq = torch.zeros(ops.MAX_STAT_ROWS, _pad64(cout), device=eng.device)
        else:
            self.ps, self.pq = eng.part_sum, eng.part_sq
        self.fsum = eng.new_stat(2 * cout_p, forward=True)     # statistics of y (also read by a PreAct block's bn1)
        if self.sliced:
            self.ps, self.pq = self.fsum[:cout_p], self.fsum[cout_p:]
            self.wf = self.wd = self.wd10 = None
        else:
            self.wf = eng.wf_view(name)
            self.wd = None if stem else eng.wd_view(name)
            self.wd10 = None if stem else eng.wd10_view(name)
        self.dw = eng.grad_flat(name + ".weight")      # storage order: OIHW, or [tap][cin][cout] (eng.tco)
        # a conv bias (VGG) is never added to y — the BatchNorm cancels it — but the running mean tracks it
        self.conv_bias = eng.param_view(name + ".bias") if (name + ".bias") in eng.p_off else None
        if bn_name is not None:
            _bn_views(self, eng, bn_name, cout_p)

    # ---- forward pieces
    def fprop(self, eng, stats=True, add_src=None):
        if self.sliced:
            C = self.cout
            xf, yf = self.x.view(-1), self.y.view(-1)
            af = add_src.view(-1) if add_src is not None else None
            for j, (ci, co) in enumerate(zip(self.g_in, self.g_out)):
                ops.conv_fprop(self.g_desc, xf[ci:], self.g_wf[j], yf[co:],
                               self.fsum[co:C] if stats else None, self.fsum[C + co:] if stats else None,
                               add_src=af[co:] if af is not None else None)
            self.stat_rows = 1
            return
        if FUSED_FINALIZE:
            ps, pq = (self.fsum[:self.cout], self.fsum[self.cout:]) if stats else (None, None)
        else:
            ps, pq = (self.ps, self.pq) if stats else (None, None)
        self.stat_rows = ops.conv_fprop(self.desc, self.x, self.wf, self.y, ps, pq, add_src=add_src)

    # ---- backward pieces
    def wgrad(self, dy):
        if self.sliced:
            xf, dyf = self.x.view(-1), dy.view(-1)
            self.eng.run_side(lambda: [ops.conv_wgrad(self.g_desc_w, xf[ci:], dyf[co:], dw, scratch=self.eng_scratch)
                                       for ci, co, dw in zip(self.g_in, self.g_out, self.g_dw)])
            return
        if self.groups > 1:
            # dense [tap][cin][cout] gradient of the block-diagonal operand, then its diagonal blocks -> dw
            cin, cout = self.cin_true, self.cout_true        # the dense matrix has the true channel counts (cin_valid)
            dense = self.eng.grouped_dense(self.k * self.k * cin * cout)

            def run():
                dense.zero_()
                ops.conv_wgrad(self.desc_w, self.x, dy, dense, scratch=self.eng_scratch)
                ops.grouped_wgrad_extract(dense, self.dw, cout, cin, self.k * self.k, self.groups)
            self.eng.run_side(run)
            return
        self.eng.run_side(lambda: ops.conv_wgrad(self.desc_w, self.x, dy, self.dw, scratch=self.eng_scratch))

    def dgrad(self, dy, dx, add_src=None):
        if self.sliced:
            dyf, dxf = dy.view(-1), dx.view(-1)
            af = add_src.view(-1) if add_src is not None else None
            for j, (ci, co) in enumerate(zip(self.g_in, self.g_out)):
                ops.conv_dgrad(self.g_desc, dyf[co:], self.g_wd[j], dxf[ci:], af[ci:] if af is not None else None)
            return
        ops.conv_dgrad(self.desc, dy, self.wd, dx, add_src)

    def dgrad_bn(self, eng, dy, dz, bn_unit, add_src=None):
        """dgrad whose result is the gradient w.r.t. relu(bn(bn_unit.y)): the epilogue applies that ReLU's mask
        and emits the BatchNorm-backward partial sums (b2c_conv_dgrad_bn). Returns the partial-row count for
        bn_unit.bn_backward_from_partial."""
        return ops.conv_dgrad_bn(self.desc, dy, self.wd, dz, bn_unit.y, bn_unit.scale, bn_unit.shift,
                                 bn_unit.bsum if FUSED_FINALIZE else eng.bwd_partial, add_src)

    def dgrad_with_shortcut(self, dy, dy_sc, sc, dx):
        """dx = dgrad(self, dy) + dgrad(sc, dy_sc): the two consumers of a down-sampling block's input
        (models/resnet.py:47-49). One launch when self is the 3x3 stride-2 conv paired with the 1x1 shortcut."""
        if self.wd10 is not None and self.k == 3 and self.stride == 2 and sc.k == 1 and sc.stride == 2:
            ops.conv_dgrad_with_shortcut(self.desc, dy, dy_sc, self.wd10, dx)
        else:
            sc.dgrad(dy_sc, dx)                       # (stride 2: zero-fill + even pixels)
            self.dgrad(dy, dx, add_src=dx)


class RootUnit(_BNOps):
    """`Root` of a DLA tree (models/dla_simple.py:44-56): relu(bn(conv1x1(cat([x0, x1], 1)))). The concatenation is
    never materialised: the forward GEMM reads both sources as the two halves of its K axis; each source's data
    gradient is a 1x1 dgrad with its own weight slab, each weight-gradient slab a 1x1 wgrad of its own source."""

    def __init__(self, eng, name, xs, C, cout, H, W):
        """xs: the 2..4 concatenated tensors ([x0, x1] in dla_simple.py; up to level + 2 = 4 in dla.py:61)."""
        N = eng.N
        self.name, self.xs, self.cin, self.cout, self.H, self.W = name + ".conv", list(xs), C, cout, H, W
        self.rows = N * H * W
        self.y = eng.new_act(N, H, W, cout)
        acc = 512 if FUSED_FINALIZE else 0
        self.desc = ops.conv_desc(N, H, W, C, cout, 1, 1, flags=acc)
        self.desc_d = ops.conv_desc(N, H, W, C, cout, 1, 1, flags=128 if CTA_PAIR == "all" else 0)
        self.desc_w = ops.conv_desc(N, H, W, C, cout, 1, 1, flags=32)
        self.ps, self.pq = eng.part_sum, eng.part_sq
        self.fsum = eng.new_stat(2 * cout, forward=True)
        self.wf, self.wd = eng.wf_view(self.name), eng.wd_view(self.name)        # [n][cout][C], [n][C][cout]
        dw = eng.grad_flat(self.name + ".weight")                                  # stored [n][C][cout]
        self.dw = [dw[t * C * cout:(t + 1) * C * cout] for t in range(len(self.xs))]
        self.eng = eng
        _bn_views(self, eng, name + ".bn", cout)

    def fprop(self, eng):
        ps, pq = (self.fsum[:self.cout], self.fsum[self.cout:]) if FUSED_FINALIZE else (self.ps, self.pq)
        if len(self.xs) == 2:
            self.stat_rows = ops.conv1x1_cat_fprop(self.desc, self.xs[0], self.xs[1], self.wf, self.y, ps, pq)
        else:
            self.stat_rows = ops.conv1x1_catn_fprop(self.desc, self.xs, self.wf, self.y, ps, pq)

    def wgrad(self, dy):
        self.eng.run_side(lambda: [ops.conv_wgrad(self.desc_w, x, dy, dw) for x, dw in zip(self.xs, self.dw)])

    def dgrad(self, dy, t, dx, add_src=None):
        ops.conv_dgrad(self.desc_d, dy, self.wd[t], dx, add_src)


class CatConvUnit(_BNOps):
    """1x1 conv (+ bias, cancelled by the BatchNorm) + BN over torch.cat(xs, 1) of tensors of different widths — the
    branch-entry convs of an Inception module (models/googlenet.py:11-15,18-21,28-31,42-44). Forward: one GEMM whose K
    axis runs through the sources' own tensor maps; each source's data gradient is a 1x1 dgrad with its own weight
    slab, each weight-gradient slab a 1x1 wgrad of its own source. With one source it is an ordinary 1x1 conv."""

    def __init__(self, eng, name, bn_name, xs, cs, cout, H, W):
        N = eng.N
        self.name, self.xs, self.cs, self.H, self.W = name, list(xs), list(cs), H, W
        self.cout_true, self.cout = cout, _pad64(cout)
        self.rows = N * H * W
        self.y = eng.new_act(N, H, W, self.cout)
        cmax = max(x.shape[-1] for x in xs)
        self.desc = ops.conv_desc(N, H, W, cmax, self.cout, 1, 1, flags=512 if FUSED_FINALIZE else 0)
        self.desc_d = [ops.conv_desc(N, H, W, x.shape[-1], self.cout, 1, 1, flags=128 if CTA_PAIR == "all" else 0)
                       for x in xs]
        self.desc_w = [ops.conv_desc(N, H, W, x.shape[-1], self.cout, 1, 1, cin_valid=c, cout_valid=cout, flags=32)
                       for x, c in zip(xs, cs)]
        self.ps, self.pq = eng.part_sum, eng.part_sq
        self.fsum = eng.new_stat(2 * self.cout, forward=True)
        self.wf = eng.wf_view(name)                              # [n][cout_p][max padded c]
        self.wd = eng.wd_slab_views(name)                        # n x [1][padded c_i][cout_p]
        dw, self.dw, row = eng.grad_flat(name + ".weight"), [], 0            # stored [sum c_i][cout]
        for c in cs:
            self.dw.append(dw[row * cout:(row + c) * cout])
            row += c
        self.conv_bias = eng.param_view(name + ".bias") if (name + ".bias") in eng.p_off else None
        self.eng = eng
        _bn_views(self, eng, bn_name, self.cout)

    def fprop(self, eng):
        ps, pq = (self.fsum[:self.cout], self.fsum[self.cout:]) if FUSED_FINALIZE else (self.ps, self.pq)
        if len(self.xs) == 1:
            self.stat_rows = ops.conv_fprop(self.desc, self.xs[0], self.wf, self.y, ps, pq)
        else:
            self.stat_rows = ops.conv1x1_catr_fprop(self.desc, self.xs, self.wf, self.y, ps, pq)

    def wgrad(self, dy):
        self.eng.run_side(lambda: [ops.conv_wgrad(d, x, dy, dw) for d, x, dw in zip(self.desc_w, self.xs, self.dw)])

    def dgrad(self, dy, t, dx, add_src=None):
        ops.conv_dgrad(self.desc_d[t], dy, self.wd[t], dx, add_src)


class DwUnit(_BNOps):
    """Depthwise 3x3 conv + BatchNorm (models/mobilenetv2.py:20-21)."""

    def __init__(self, eng, name, bn_name, x, C, stride, H, W, src_unit=None):
        """src_unit: the conv + BatchNorm unit whose relu(bn(y)) is this conv's input, when that activation is not
        materialised (x is None): the kernels read src_unit.y and apply src_unit.scale / shift + ReLU on load."""
        N = eng.N
        self.src = src_unit
        self.name, self.c_true, self.stride, self.H, self.W = name, C, stride, H, W
        self.cout = _pad64(C)
        self.Ho, self.Wo = H // stride, W // stride
        self.rows = N * self.Ho * self.Wo
        self.x = x
        self.y = eng.new_act(N, self.Ho, self.Wo, self.cout)
        self.w = eng.param_view(name + ".weight")
        self.dw = eng.grad_view(name + ".weight")
        self.eng = eng
        self.ps, self.pq = eng.part_sum, eng.part_sq
        self.conv_bias = None
        _bn_views(self, eng, bn_name, self.cout)

    def fprop(self, eng):
        if self.src is not None:
            u = self.src
            if FUSED_FINALIZE:
                self.stat_rows = ops.dwconv_fprop_bnrelu(u.y, u.scale, u.shift, self.w, self.y, self.c_true,
                                                         self.fsum[:self.cout], self.fsum[self.cout:], accumulate=True)
            else:
                self.stat_rows = ops.dwconv_fprop_bnrelu(u.y, u.scale, u.shift, self.w, self.y, self.c_true,
                                                         eng.part_sum, eng.part_sq)
        elif FUSED_FINALIZE:
            ops.dwconv_fprop_acc(self.x, self.w, self.y, self.c_true, self.stride, self.fsum[:self.cout],
                                 self.fsum[self.cout:])
            self.stat_rows = 1
        else:
            self.stat_rows = ops.dwconv_fprop(self.x, self.w, self.y, self.c_true, self.stride, eng.part_sum,
                                              eng.part_sq)

    def wgrad(self, dy):
        if self.src is not None:
            u = self.src
            self.eng.run_side(lambda: ops.dwconv_wgrad_bnrelu(u.y, u.scale, u.shift, dy, self.dw, self.c_true))
        else:
            self.eng.run_side(lambda: ops.dwconv_wgrad(self.x, dy, self.dw, self.c_true, self.stride))

    def dgrad(self, dy, dx):
        ops.dwconv_dgrad(dy, self.w, dx, self.c_true, self.stride)


class Engine:
    """Static execution plan for one (network, batch size) pair on one GPU."""

    def __init__(self, spec, batch, device="cuda", share_from=None):
        """share_from: another Engine of the same spec whose parameter / buffer arenas are reused (e.g. an
        eval-batch plan next to the training plan)."""
        self.spec, self.N, self.device = spec, batch, torch.device(device)
        self._layout_params(share_from)
        self._acts = []
        self._side, self._side_dirty = None, False
        f32 = dict(device=self.device, dtype=torch.float32)
        stat_rows = ops.MAX_DW_STAT_ROWS if spec["family"] in ("mobilenetv2", "mobilenet") else ops.MAX_STAT_ROWS
        self.part_sum = torch.zeros(stat_rows, 2048, **f32)
        self.part_sq = torch.zeros(stat_rows, 2048, **f32)
        self.bwd_partial = torch.zeros(148 * 4 * 2 * 2048, **f32)
        self.coef_a, self.coef_b, self.coef_c = (torch.zeros(2048, **f32) for _ in range(3))
        self.wgrad_scratch = share_from.wgrad_scratch if share_from is not None else \
            torch.zeros(ops.WGRAD_SCRATCH_FLOATS, **f32)
        # BatchNorm statistic accumulators (FUSED_FINALIZE): [sum | sumsq] per BatchNorm for the forward pass, zeroed by
        # forward(); [sum dz | sum dz*y] for the backward pass, zeroed by zero_grad().
        nstat = sum(2 * _pad64(sh[0]) for name, sh in spec["buffers"].items() if name.endswith("running_mean"))
        nconv = sum(2 * _pad64(sh[0]) for name, sh in spec["params"].items() if len(sh) == 4)
        self._stat_f = torch.zeros(nstat + nconv + 64, **f32)
        self._stat_b = torch.zeros(nstat + 64, **f32)
        self._stat_used = [0, 0]
        self.fwd_ops, self.bwd_ops = [], []
        self.stem_taps = 9
        self.zero_stats_each_forward = False
        if spec["family"] == "resnet":
            self._build_resnet()
        elif spec["family"] == "vgg":
            self._build_vgg()
        elif spec["family"] == "preact":
            self._build_preact()
        elif spec["family"] == "mobilenetv2":
            self._build_mobilenetv2()
        elif spec["family"] == "bottleneck":
            self._build_bottleneck()
        elif spec["family"] == "dla":
            self._build_dla()
        elif spec["family"] == "mobilenet":
            self._build_mobilenet()
        elif spec["family"] == "senet":
            self._build_senet()
        elif spec["family"] == "googlenet":
            self._build_googlenet()
        else:
            raise NotImplementedError(spec["family"])

    # ------------------------------------------------------------------ arenas
    def _layout_params(self, share_from=None):
        spec = self.spec
        self.p_off, self.n_param = param_offsets(spec)
        self.b_off, off = OrderedDict(), 0
        for name, shape in spec["buffers"].items():
            self.b_off[name] = (off, shape[0])
            off = _align(off + _pad64(shape[0]), 4)          # padded like the per-channel parameters
        f32 = dict(device=self.device, dtype=torch.float32)
        if share_from is not None:
            self.P, self.G, self.M, self.B = share_from.P, share_from.G, share_from.M, share_from.B
        else:
            self.P = torch.zeros(self.n_param, **f32)      # master parameters (reference layouts: OIHW etc.)
            self.G = torch.zeros(self.n_param, **f32)      # gradients, same layout (= the reference's .grad)
            self.M = torch.zeros(self.n_param, **f32)      # SGD momentum buffers
            self.B = torch.zeros(max(off, 4), **f32)       # BN running statistics
        # bf16 GEMM operand copies of the conv weights
        self.wf_off, self.wd_off, entries = {}, {}, []
        # conv weights stored [tap][cin][cout] in the P / G / M arenas (see _arena_view): every dense conv except the
        # 3-channel stem (its wgrad addresses the im2col'ed k axis) with channel counts the vector path can take
        grouped = spec.get("groups", {})
        not_conv = set(spec.get("depthwise", ())) | set(spec.get("se_fc", ()))     # 4-D weights the GEMM kernels never see
        self.tco = set() if not TCO_LAYOUT else {
            name for name, shape in spec["params"].items()
            if len(shape) == 4 and name not in not_conv and name not in grouped and shape[1] != 3
            and shape[0] % 4 == 0 and shape[1] % 2 == 0}
        self.wd_pair = {}      # 3x3 stride-2 conv1 -> its block's 1x1 shortcut conv, stored right behind it in WD
        self.wd_slabs = {}     # 1x1 conv over a ragged concatenation -> [(offset, shape)] of its per-source dgrad operands
        self.group_ops = {}    # grouped conv run group by group -> [(wf offset, wd offset, wf shape, wd shape)]
        wf_n = wd_n = 0
        convs = [(name, shape) for name, shape in spec["params"].items() if len(shape) == 4 and name not in not_conv]
        # A down-sampling block's `conv1` (3x3) and `shortcut.0` (1x1) read the same input and have the same
        # channel counts: their dgrad operands sit back to back so ONE 10-slab matrix serves the fused dgrad
        # (b2c_conv_dgrad_with_shortcut).
        shapes = dict(convs)
        order, placed = [], set()
        for name, shape in convs:
            if name in placed:
                continue
            order.append(name); placed.add(name)
            if name.endswith(".conv1.weight") and shape[2] == 3:
                sc = name[:-len("conv1.weight")] + "shortcut.0.weight"
                if sc in shapes and shapes[sc][:2] == shape[:2] and shapes[sc][2] == 1 and sc not in placed:
                    order.append(sc); placed.add(sc)
                    self.wd_pair[name[:-len(".weight")]] = sc[:-len(".weight")]
        for name in order:
            shape = shapes[name]
            cname = name[:-len(".weight")]
            cout, cin, k, _ = shape
            cout_p, cin_p = _pad64(cout), _pad64(cin)
            poff = self.p_off[name][0]
            if name in spec.get("cat_src", ()):
                # 1x1 conv over a concatenation of tensors of different widths (models/googlenet.py:52-57): the master
                # [1][sum c_i][cout] is one slab of c_i rows per source. fprop operand: [n][cout][max padded c_i], slab i
                # filled in its first c_i columns (b2c_conv1x1_catr_fprop); dgrad operands: one [padded c_i][cout] each
                cs = spec["cat_src"][name]
                assert name in self.tco and k == 1 and sum(cs) == cin, name
                cmax_p = max(_pad64(c) for c in cs)
                self.wf_off[cname] = (wf_n, (len(cs), cout_p, cmax_p))
                self.wd_slabs[cname] = []
                row = 0
                for i, c in enumerate(cs):
                    self.wd_slabs[cname].append((wd_n, (1, _pad64(c), cout_p)))
                    entries.append(dict(src_off=poff + row * cout, wf_off=wf_n + i * cout_p * cmax_p, wd_off=wd_n,
                                        cout=cout, cin=c, taps=1, cin_pad=cmax_p, cout_pad=cout_p, kflat=2))
                    wd_n += _pad64(c) * cout_p
                    row += c
                wf_n += len(cs) * cout_p * cmax_p
            elif name in spec.get("catn", ()):
                # Root conv over cat(xs) (models/dla_simple.py:52-54, models/dla.py:47-50): the [1][n*C][cout] master IS
                # [n][C][cout], so the operands are those of an n-tap conv with C input channels
                # (b2c_conv1x1_cat_fprop / b2c_conv1x1_catn_fprop)
                ns = spec["catn"][name]
                assert name in self.tco and k == 1 and cin % (64 * ns) == 0, name
                cin = cin // ns
                cin_p = _pad64(cin)
                self.wf_off[cname] = (wf_n, (ns, cout_p, cin_p))
                self.wd_off[cname] = (wd_n, (ns, cin_p, cout_p))
                entries.append(dict(src_off=poff, wf_off=wf_n, wd_off=wd_n, cout=cout, cin=cin, taps=ns,
                                    cin_pad=cin_p, cout_pad=cout_p, kflat=2))
                wf_n += ns * cout_p * cin_p
                wd_n += ns * cout_p * cin_p
            elif name in grouped and GROUP_SLICES and cin % 64 == 0 and (cout // grouped[name]) % 64 == 0:
                # nn.Conv2d(groups=g) with group widths on the 64-channel grid: one dense cin/g -> cout/g conv per
                # group on a channel slice of the activations (B2cConvDesc.x_ld / y_ld); its operands are the group's
                # rows of the [cout][cin/g][kh][kw] master, which are a dense OIHW weight of their own
                g = grouped[name]
                cin_g, cout_g = cin, cout // g
                self.group_ops[cname] = []
                for j in range(g):
                    self.group_ops[cname].append((wf_n, wd_n, (k * k, cout_g, cin_g), (k * k, cin_g, cout_g)))
                    entries.append(dict(src_off=poff + j * cout_g * cin_g * k * k, wf_off=wf_n, wd_off=wd_n,
                                        cout=cout_g, cin=cin_g, taps=k * k, cin_pad=cin_g, cout_pad=cout_g, kflat=0))
                    wf_n += cout_g * cin_g * k * k
                    wd_n += cout_g * cin_g * k * k
            elif name in grouped:
                # nn.Conv2d(groups=g): master [cout][cin/g][kh][kw]; the operands are those of the dense cin -> cout
                # conv, filled block-diagonally by one entry per group (off-diagonal blocks stay zero)
                g = grouped[name]
                cin_g, cout_g, cin = cin, cout // g, cin * g
                cin_p = _pad64(cin)          # (padded channels: rows / columns beyond the last group stay zero too)
                self.wf_off[cname] = (wf_n, (k * k, cout_p, cin_p))
                self.wd_off[cname] = (wd_n, (k * k, cin_p, cout_p))
                for j in range(g):
                    entries.append(dict(src_off=poff + j * cout_g * cin_g * k * k,
                                        wf_off=wf_n + j * cout_g * cin_p + j * cin_g,
                                        wd_off=wd_n + j * cin_g * cout_p + j * cout_g,
                                        cout=cout_g, cin=cin_g, taps=k * k, cin_pad=cin_p, cout_pad=cout_p, kflat=0))
                wf_n += cout_p * cin_p * k * k
                wd_n += cout_p * cin_p * k * k
            elif cin == 3:      # stem: [cout_pad][64], k = tap*3 + c
                self.wf_off[cname] = (wf_n, (1, cout_p, 64))
                entries.append(dict(src_off=poff, wf_off=wf_n, wd_off=-1, cout=cout, cin=cin, taps=k * k,
                                    cin_pad=64, kflat=1))
                wf_n += cout_p * 64
            else:
                self.wf_off[cname] = (wf_n, (k * k, cout_p, cin_p))
                self.wd_off[cname] = (wd_n, (k * k, cin_p, cout_p))
                entries.append(dict(src_off=poff, wf_off=wf_n, wd_off=wd_n, cout=cout, cin=cin, taps=k * k,
                                    cin_pad=cin_p, cout_pad=cout_p, kflat=2 if name in self.tco else 0))
                wf_n += cout_p * cin_p * k * k
                wd_n += cout_p * cin_p * k * k
        bf = dict(device=self.device, dtype=torch.bfloat16)
        self.repack_entries = len(entries)
        if share_from is not None:
            self.WF, self.WD, self.repack_table = share_from.WF, share_from.WD, share_from.repack_table
        else:
            self.WF = torch.zeros(_align(wf_n, 8), **bf)
            self.WD = torch.zeros(_align(max(wd_n, 8), 8), **bf)
            self.repack_table = ops.make_repack_table(entries, self.device)

    def wd10_view(self, cname):
        """[10][cin][cout] dgrad operand of a paired conv1 + shortcut (None if `cname` has no pair)."""
        if cname not in self.wd_pair:
            return None
        off, (taps, cin_p, cout_p) = self.wd_off[cname]
        off2, _ = self.wd_off[self.wd_pair[cname]]
        assert taps == 9 and off2 == off + 9 * cin_p * cout_p
        return self.WD[off:off + 10 * cin_p * cout_p].view(10, cin_p, cout_p)

    def _arena_view(self, arena, name):
        """The reference-shaped (OIHW, ...) tensor `name` inside an arena. Conv weights in self.tco are STORED
        [kh][kw][cin][cout] (the order the wgrad kernels reduce into with coalesced vector atomics, and the dgrad
        operand's own order); the view handed out is that storage permuted back to OIHW, so every consumer of
        state_dict / .grad semantics (load_params, named_params, accelerate(), optimizers, checkpoints) sees the
        reference's shape and element meaning. Element-wise arena kernels (SGD, all-reduce) do not care."""
        off, n, shape = self.p_off[name]
        if name in self.tco:
            co, ci, kh, kw = shape
            return arena[off:off + n].view(kh, kw, ci, co).permute(3, 2, 0, 1)
        return arena[off:off + n].view(shape)

    def param_view(self, name):
        return self._arena_view(self.P, name)

    def grad_view(self, name):
        return self._arena_view(self.G, name)

    def grad_flat(self, name):
        """Contiguous storage of a parameter's gradient (what the kernels write)."""
        off, n, _ = self.p_off[name]
        return self.G[off:off + n]

    def buffer_view(self, name):
        off, n = self.b_off[name]
        return self.B[off:off + n]

    # padded-width views of per-channel vectors, for the kernels (the tail beyond the true length stays zero)
    def param_vec(self, name, C):
        off = self.p_off[name][0]
        return self.P[off:off + C]

    def grad_vec(self, name, C):
        off = self.p_off[name][0]
        return self.G[off:off + C]

    def buffer_vec(self, name, C):
        off = self.b_off[name][0]
        return self.B[off:off + C]

    def wf_view(self, cname):
        off, shape = self.wf_off[cname]
        n = shape[0] * shape[1] * shape[2]
        return self.WF[off:off + n].view(shape)

    def wd_slab_views(self, cname):
        return [self.WD[off:off + sh[0] * sh[1] * sh[2]].view(sh) for off, sh in self.wd_slabs[cname]]

    def wd_view(self, cname):
        off, shape = self.wd_off[cname]
        n = shape[0] * shape[1] * shape[2]
        return self.WD[off:off + n].view(shape)

    def new_act(self, n, h, w, c):
        t = torch.zeros(n, h, w, c, device=self.device, dtype=torch.bfloat16)
        self._acts.append(t)
        return t

    def new_f32(self, *shape):
        return torch.zeros(*shape, device=self.device, dtype=torch.float32)

    def grouped_dense(self, n):
        """Dense weight-gradient matrix of a grouped conv (ConvUnit.wgrad): one buffer, sized for the largest, shared
        by all grouped convs of the plan — they use it one after the other on the side stream."""
        if getattr(self, "_grouped_dense", None) is None or self._grouped_dense.numel() < n:   # (plan construction)
            self._grouped_dense = torch.zeros(n, device=self.device, dtype=torch.float32)
        return self._grouped_dense[:n]

    def new_stat(self, n, forward):
        """n floats of the forward / backward statistic accumulator arena (16-byte aligned)."""
        arena, k = (self._stat_f, 0) if forward else (self._stat_b, 1)
        off = self._stat_used[k]
        if off + n > arena.numel():
            raise RuntimeError("statistic arena too small for this plan")
        self._stat_used[k] = _align(off + n, 4)
        return arena[off:off + n]

    # ------------------------------------------------------------------ parameters in / out
    def load_params(self, named_params, named_buffers=None):
        """Copy reference-layout tensors (name -> tensor, fp32) into the arenas and refresh the bf16 copies."""
        for name, (off, n, shape) in self.p_off.items():
            self.param_view(name).copy_(named_params[name].detach().reshape(shape).to(self.device, torch.float32))
        if named_buffers is not None:
            for name, (off, n) in self.b_off.items():
                self.B[off:off + n].copy_(named_buffers[name].detach().to(self.device, torch.float32))
        self.repack()

    def init_bn_buffers(self):
        for name, (off, n) in self.b_off.items():
            self.B[off:off + n].fill_(1.0 if name.endswith("running_var") else 0.0)

    def repack(self):
        ops.repack_weights(self.P, self.WF, self.WD, self.repack_table, self.repack_entries)

    def named_grads(self):
        return OrderedDict((name, self.grad_view(name)) for name in self.p_off)

    def named_params(self):
        return OrderedDict((name, self.param_view(name)) for name in self.p_off)

    def named_buffers(self):
        return OrderedDict((name, self.buffer_view(name)) for name in self.b_off)

    # ------------------------------------------------------------------ ResNet plan
    def _build_resnet(self):
        spec, N = self.spec, self.N
        self.x0 = self.new_act(N, 32, 32, 64)            # stem im2col operand (k = 27 of 64)
        stem = ConvUnit(self, "conv1", "bn1", self.x0, 3, 64, 3, 1, 32, 32, stem=True)
        a0 = self.new_act(N, 32, 32, 64)
        self.units = [stem]
        F, Bk = self.fwd_ops, []

        def fwd_stem(training):
            stem.fprop(self, True)
            stem.bn_apply(self, training, a0)
        F.append(fwd_stem)

        x, H, W = a0, 32, 32
        blocks = []
        for b in spec["blocks"]:
            name, cin, cout, s = b["name"], b["cin"], b["cout"], b["stride"]
            c1 = ConvUnit(self, name + ".conv1", name + ".bn1", x, cin, cout, 3, s, H, W)
            Ho, Wo = H // s, W // s
            a1 = self.new_act(N, Ho, Wo, cout)
            c2 = ConvUnit(self, name + ".conv2", name + ".bn2", a1, cout, cout, 3, 1, Ho, Wo)
            cs = None
            if b["shortcut"]:
                cs = ConvUnit(self, name + ".shortcut.0", name + ".shortcut.1", x, cin, cout, 1, s, H, W,
                              private_stats=True)
            out = self.new_act(N, Ho, Wo, cout)
            bits = torch.zeros(out.numel() // 8, dtype=torch.uint8, device=self.device) if RELU_BITS else None
            self.units += [c1, c2] + ([cs] if cs else [])
            blocks.append((c1, a1, c2, cs, out, x, bits))

            def fwd_block(training, c1=c1, a1=a1, c2=c2, cs=cs, out=out, x=x, bits=bits):
                if cs is not None:          # the shortcut branch (own statistics scratch) overlaps conv1 -> conv2
                    self.run_side(lambda: (cs.fprop(self), None if (FUSED_FINALIZE and training) else cs.bn_stats(self, training)))
                c1.fprop(self)
                c1.bn_apply(self, training, a1)
                c2.fprop(self)
                if cs is not None:
                    self.join_side()
                    c2.bn_apply(self, training, out, res_unit=cs, relu_bits=bits)
                else:
                    c2.bn_apply(self, training, out, res=x, relu_bits=bits)
            F.append(fwd_block)
            x, H, W = out, Ho, Wo

        # head
        self.feat_act, self.feat_hw, self.feat_c = x, H * W, spec["feat"]
        K = spec["num_classes"]
        self.lin_w, self.lin_b = self.param_view("linear.weight"), self.param_view("linear.bias")
        self.pooled = self.new_f32(N, self.feat_c)
        self.logits = self.new_f32(N, K)
        self.dlogits = self.new_f32(N, K)

        def fwd_head(training):
            ops.head_fwd(self.feat_act, N, self.feat_hw, self.feat_c, self.lin_w, self.lin_b, K, self.pooled,
                         self.logits)
        F.append(fwd_head)

        # ---- backward: gradient buffers. g_* hold d(block output); two ping-pong sets per resolution are
        # enough because a block's input gradient is complete before the previous block consumes it.
        def bwd_head():
            ops.head_bwd(self.dlogits, self.pooled, self.lin_w, N, self.feat_hw, self.feat_c, K, dout_last,
                         self.grad_view("linear.weight"), self.grad_view("linear.bias"))

        dout_last = self.new_act(*self.feat_act.shape)
        Bk.append(bwd_head)
        dout = dout_last
        self.stem_unit, self.a0 = stem, a0
        self.block_bufs = []      # backward order: tensors of each block, for tests / introspection
        for (c1, a1, c2, cs, out, xin, bits) in reversed(blocks):
            dy2 = self.new_act(*c2.y.shape)
            da1 = self.new_act(*a1.shape)
            dy1 = self.new_act(*c1.y.shape)
            dys = self.new_act(*cs.y.shape) if cs is not None else None
            dx = self.new_act(*xin.shape)

            def bwd_block(c1=c1, a1=a1, c2=c2, cs=cs, out=out, dout=dout, dy2=dy2, da1=da1, dy1=dy1, dys=dys,
                          dx=dx, bits=bits):
                # out = relu(bn2(y2) + shortcut): mask in place -> dz, then BN2 backward
                c2.bn_backward(self, dout, out, dout, dy2, relu_bits=bits)
                if cs is not None:
                    cs.bn_backward(self, dout, None, None, dys)
                c2.wgrad(dy2)
                # a1 = relu(bn1(y1)): its mask and BN-backward sums come out of conv2's dgrad epilogue
                if FUSE_DGRAD_BN:
                    prow = c2.dgrad_bn(self, dy2, da1, c1)
                    c1.bn_backward_from_partial(self, da1, dy1, prow)
                else:
                    c2.dgrad(dy2, da1)
                    c1.bn_backward(self, da1, a1, da1, dy1, relu_from_y=True)
                c1.wgrad(dy1)
                if cs is not None:
                    cs.wgrad(dys)
                    c1.dgrad_with_shortcut(dy1, dys, cs, dx)
                else:
                    c1.dgrad(dy1, dx, add_src=dout)       # identity shortcut: + dz
            Bk.append(bwd_block)
            self.block_bufs.append(dict(c1=c1, a1=a1, c2=c2, cs=cs, out=out, xin=xin, dout=dout, dx=dx, bits=bits))
            dout = dx

        dy0 = self.new_act(*stem.y.shape)

        def bwd_stem(dout=dout, dy0=dy0):
            stem.bn_backward(self, dout, a0, dout, dy0, relu_from_y=True)
            stem.wgrad(dy0)
        Bk.append(bwd_stem)
        self.bwd_ops = Bk

    # ------------------------------------------------------------------ SimpleDLA plan (models/dla_simple.py)
    def _build_dla(self):
        """models/dla_simple.py:108-116: base -> layer1 -> layer2 (conv3x3 + BN + ReLU each) -> four Trees ->
        avg_pool2d(4) -> Linear. Tree.forward (:72-76): out1 = left(x); out2 = right(out1); root([out1, out2]).
        Backward of a tree: root (BN/ReLU backward, both weight-gradient slabs, out2's data gradient), the right
        sub-tree, then out1's gradient = root slab 0 + what the right sub-tree returned, then the left sub-tree."""
        spec, N = self.spec, self.N
        self.x0 = self.new_act(N, 32, 32, 64)
        self.units = []
        F = self.fwd_ops
        eng = self

        class Block:                     # BasicBlock, models/dla_simple.py:16-41 (= models/resnet.py:16-51)
            def __init__(b, d, x, H, W):
                name, cin, cout, s = d["name"], d["cin"], d["cout"], d["stride"]
                b.x, b.H, b.W = x, H // s, W // s
                b.c1 = ConvUnit(eng, name + ".conv1", name + ".bn1", x, cin, cout, 3, s, H, W)
                b.a1 = eng.new_act(N, b.H, b.W, _pad64(cout))
                b.c2 = ConvUnit(eng, name + ".conv2", name + ".bn2", b.a1, cout, cout, 3, 1, b.H, b.W)
                b.cs = ConvUnit(eng, name + ".shortcut.0", name + ".shortcut.1", x, cin, cout, 1, s, H, W,
                                private_stats=True) if d["shortcut"] else None
                b.out = eng.new_act(N, b.H, b.W, _pad64(cout))
                b.bits = torch.zeros(b.out.numel() // 8, dtype=torch.uint8, device=eng.device)
                eng.units += [b.c1, b.c2] + ([b.cs] if b.cs else [])

            def fwd(b, training):
                c1, c2, cs = b.c1, b.c2, b.cs
                if cs is not None:
                    eng.run_side(lambda: (cs.fprop(eng), None if (FUSED_FINALIZE and training) else cs.bn_stats(eng, training)))
                c1.fprop(eng)
                c1.bn_apply(eng, training, b.a1)
                c2.fprop(eng)
                if cs is not None:
                    eng.join_side()
                    c2.bn_apply(eng, training, b.out, res_unit=cs, relu_bits=b.bits)
                else:
                    c2.bn_apply(eng, training, b.out, res=b.x, relu_bits=b.bits)

            def plan_bwd(b, dout, ops_list, acc=None):
                """Appends this block's backward to ops_list; dout = gradient buffer of b.out. Returns d(b.x).
                acc: a buffer that already holds another consumer's gradient of b.x — this block's is added to it in
                place (conv-shortcut blocks only: each of the two data-gradient kernels accumulates)."""
                c1, c2, cs = b.c1, b.c2, b.cs
                dy2, da1, dy1 = (eng.new_act(*t.shape) for t in (c2.y, b.a1, c1.y))
                dys = eng.new_act(*cs.y.shape) if cs is not None else None
                if acc is not None and cs is None:
                    raise NotImplementedError("accumulating input gradient of an identity-shortcut block")
                dx = eng.new_act(*b.x.shape) if acc is None else acc

                def run():
                    c2.bn_backward(eng, dout, b.out, dout, dy2, relu_bits=b.bits)
                    if cs is not None:
                        cs.bn_backward(eng, dout, None, None, dys)
                    c2.wgrad(dy2)
                    c1.bn_backward_from_partial(eng, da1, dy1, c2.dgrad_bn(eng, dy2, da1, c1))
                    c1.wgrad(dy1)
                    if cs is not None:
                        cs.wgrad(dys)
                        if acc is not None:
                            cs.dgrad(dys, dx, add_src=dx)
                            c1.dgrad(dy1, dx, add_src=dx)
                        else:
                            c1.dgrad_with_shortcut(dy1, dys, cs, dx)
                    else:
                        c1.dgrad(dy1, dx, add_src=dout)
                ops_list.append(run)
                return dx

        class Tree:
            def __init__(t, d, x, H, W):
                sub = Tree if d["left"]["kind"] == "tree" else Block
                t.left = sub(d["left"], x, H, W)
                t.right = sub(d["right"], t.left.out, t.left.H, t.left.W)
                t.H, t.W = t.left.H, t.left.W
                t.root = RootUnit(eng, d["name"] + ".root", [t.left.out, t.right.out], d["cout"], d["cout"], t.H, t.W)
                t.out = eng.new_act(N, t.H, t.W, d["cout"])
                eng.units.append(t.root)

            def fwd(t, training):
                t.left.fwd(training)
                t.right.fwd(training)
                t.root.fprop(eng)
                t.root.bn_apply(eng, training, t.out)

            def plan_bwd(t, dout, ops_list, acc=None):
                root = t.root
                dyr = eng.new_act(*root.y.shape)
                g2 = eng.new_act(*t.right.out.shape)

                def run_root():
                    root.bn_backward(eng, dout, t.out, dout, dyr, relu_from_y=True)
                    root.wgrad(dyr)
                    root.dgrad(dyr, 1, g2)
                ops_list.append(run_root)
                g1 = t.right.plan_bwd(g2, ops_list)                  # gradient of out1 through the right sub-tree
                ops_list.append(lambda: root.dgrad(dyr, 0, g1, add_src=g1))      # + through the root's first slab
                return t.left.plan_bwd(g1, ops_list, acc) if acc is not None else t.left.plan_bwd(g1, ops_list)

        class Tree2:
            """Level-2 tree of models/dla.py:57-84: prev_root(x) and level_1(x) both read the tree's input; left_node
            reads level_1's output, right_node left_node's; the root concatenates all four. Backward: root, then
            right_node, left_node, level_1 — each producer's gradient = its root slab + what its consumer returned —
            then prev_root, whose input gradient is added to level_1's."""
            def __init__(t, d, x, H, W):
                t.prev = Block(d["prev_root"], x, H, W)
                t.l1 = Tree(d["level_1"], x, H, W)
                t.H, t.W = t.l1.H, t.l1.W
                t.left = Block(d["left"], t.l1.out, t.H, t.W)
                t.right = Block(d["right"], t.left.out, t.H, t.W)
                t.root = RootUnit(eng, d["name"] + ".root", [t.prev.out, t.l1.out, t.left.out, t.right.out], d["cout"],
                                  d["cout"], t.H, t.W)
                t.out = eng.new_act(N, t.H, t.W, d["cout"])
                eng.units.append(t.root)

            def fwd(t, training):
                t.prev.fwd(training)
                t.l1.fwd(training)
                t.left.fwd(training)
                t.right.fwd(training)
                t.root.fprop(eng)
                t.root.bn_apply(eng, training, t.out)

            def plan_bwd(t, dout, ops_list):
                root = t.root
                dyr = eng.new_act(*root.y.shape)
                g_right = eng.new_act(*t.right.out.shape)
                g_prev = eng.new_act(*t.prev.out.shape)

                def run_root():
                    root.bn_backward(eng, dout, t.out, dout, dyr, relu_from_y=True)
                    root.wgrad(dyr)
                    root.dgrad(dyr, 3, g_right)
                    root.dgrad(dyr, 0, g_prev)
                ops_list.append(run_root)
                g_left = t.right.plan_bwd(g_right, ops_list)
                ops_list.append(lambda: root.dgrad(dyr, 2, g_left, add_src=g_left))
                g_l1 = t.left.plan_bwd(g_left, ops_list)
                ops_list.append(lambda: root.dgrad(dyr, 1, g_l1, add_src=g_l1))
                dx = t.prev.plan_bwd(g_prev, ops_list)               # d(x) through prev_root ...
                return t.l1.plan_bwd(g_l1, ops_list, acc=dx)         # ... + through level_1, accumulated in place

        x, H, W = self.x0, 32, 32
        plain = []
        for i, (name, cin, cout) in enumerate(spec["plain"]):
            u = ConvUnit(self, name + ".0", name + ".1", x, cin, cout, 3, 1, 32, 32, stem=(i == 0))
            act = self.new_act(N, 32, 32, u.cout)
            self.units.append(u)
            plain.append((u, act))
            F.append(lambda training, u=u, act=act: (u.fprop(self), u.bn_apply(self, training, act)))
            x = act
        trees = []
        for d in spec["trees"]:
            t = (Tree2 if d["kind"] == "tree2" else Tree)(d, x, H, W)
            trees.append(t)
            F.append(t.fwd)
            x, H, W = t.out, t.H, t.W
        fwd_head, bwd_head, dout = self._build_head(x, H, W, "linear.weight", "linear.bias")
        F.append(fwd_head)
        Bk = [bwd_head]
        for t in reversed(trees):
            sub = []
            dout = t.plan_bwd(dout, sub)
            Bk.append(lambda sub=sub: [f() for f in sub])
        for i in range(len(plain) - 1, -1, -1):
            u, act = plain[i]
            dy = self.new_act(*u.y.shape)
            dx = self.new_act(*u.x.shape) if i > 0 else None

            def bwd_plain(u=u, act=act, dout=dout, dy=dy, dx=dx):
                u.bn_backward(self, dout, act, dout, dy, relu_from_y=True)
                u.wgrad(dy)
                if dx is not None:
                    u.dgrad(dy, dx)
            Bk.append(bwd_plain)
            dout = dx
        self.bwd_ops = Bk

    # ------------------------------------------------------------------ shared head (pool + linear)
    def _build_head(self, x, H, W, wname, bname):
        spec, N = self.spec, self.N
        self.feat_act, self.feat_hw, self.feat_c = x, H * W, spec["feat"]
        K = spec["num_classes"]
        self.lin_w, self.lin_b = self.param_view(wname), self.param_view(bname)
        self.pooled = self.new_f32(N, self.feat_c)
        self.logits = self.new_f32(N, K)
        self.dlogits = self.new_f32(N, K)
        dout_last = self.new_act(*x.shape)

        ld = x.shape[-1]                     # channel pitch of the last activation (padded to 64: RegNetX ends on 368)

        def fwd_head(training):
            ops.head_fwd(self.feat_act, N, self.feat_hw, self.feat_c, self.lin_w, self.lin_b, K, self.pooled,
                         self.logits, act_ld=ld)

        def bwd_head():
            ops.head_bwd(self.dlogits, self.pooled, self.lin_w, N, self.feat_hw, self.feat_c, K, dout_last,
                         self.grad_view(wname), self.grad_view(bname), act_ld=ld)
        return fwd_head, bwd_head, dout_last

    # ------------------------------------------------------------------ VGG plan (models/vgg.py:20-38)
    def _build_vgg(self):
        spec, N = self.spec, self.N
        self.x0 = self.new_act(N, 32, 32, 64)
        self.units = []
        F, stages = self.fwd_ops, []
        x, H, W = self.x0, 32, 32
        layers = spec["layers"]
        for li, layer in enumerate(layers):
            if layer[0] != "conv":
                continue
            _, cname, bname, cin, cout = layer
            u = ConvUnit(self, cname, bname, x, cin, cout, 3, 1, H, W, stem=(cin == 3))
            act = self.new_act(N, H, W, cout)
            pooled = None
            if li + 1 < len(layers) and layers[li + 1][0] == "pool":
                pooled = self.new_act(N, H // 2, W // 2, cout)
            self.units.append(u)
            stages.append((u, act, pooled, x))

            def fwd(training, u=u, act=act, pooled=pooled):
                u.fprop(self)
                u.bn_apply(self, training, act)
                if pooled is not None:
                    ops.maxpool2x2_fwd(act, pooled)
            F.append(fwd)
            if pooled is not None:
                x, H, W = pooled, H // 2, W // 2
            else:
                x = act
        fwd_head, bwd_head, dout = self._build_head(x, H, W, "classifier.weight", "classifier.bias")
        F.append(fwd_head)
        Bk = [bwd_head]
        for si in range(len(stages) - 1, -1, -1):
            u, act, pooled, xin = stages[si]
            dact = self.new_act(*act.shape) if pooled is not None else None
            dy = self.new_act(*u.y.shape)
            dx = self.new_act(*xin.shape) if si > 0 else None

            # the unit below (si-1) takes its BN-backward sums from this unit's dgrad epilogue when no max-pool
            # sits between them; `fused_in` says whether the unit above did that for us
            prev = stages[si - 1] if si > 0 else None
            fuse_out = FUSE_DGRAD_BN and prev is not None and prev[2] is None
            fused_in = FUSE_DGRAD_BN and si + 1 < len(stages) and pooled is None
            prow_box = self._vgg_prow = getattr(self, "_vgg_prow", {})

            def bwd(u=u, act=act, pooled=pooled, dout=dout, dact=dact, dy=dy, dx=dx, si=si, prev=prev,
                    fuse_out=fuse_out, fused_in=fused_in):
                g = dout
                if pooled is not None:
                    ops.maxpool2x2_bwd(act, pooled, dout, dact)
                    g = dact
                if fused_in:
                    u.bn_backward_from_partial(self, g, dy, prow_box[si])
                else:
                    u.bn_backward(self, g, act, g, dy, relu_from_y=True)
                u.wgrad(dy)
                if dx is not None:
                    if fuse_out:
                        prow_box[si - 1] = u.dgrad_bn(self, dy, dx, prev[0])
                    else:
                        u.dgrad(dy, dx)
            Bk.append(bwd)
            dout = dx
        self.bwd_ops = Bk

    # ------------------------------------------------------------------ PreActResNet plan
    def _build_preact(self):
        """models/preact_resnet.py:28-34: out = relu(bn1(x)); sc = shortcut(out) | x; out = conv1(out);
        out = conv2(relu(bn2(out))); out += sc. BatchNorm sits BEFORE each conv, so a BN's statistics come
        from the epilogue of whichever conv produced its input (conv2's epilogue includes the residual add)."""
        spec, N = self.spec, self.N
        self.x0 = self.new_act(N, 32, 32, 64)
        stem = ConvUnit(self, "conv1", None, self.x0, 3, 64, 3, 1, 32, 32, stem=True)
        self.units = [stem]
        F, blocks = self.fwd_ops, []

        class PreBN(_BNOps):
            """BatchNorm applied to tensor `y` produced by `producer` (whose epilogue emitted the statistics)."""
            def __init__(s, eng, bn_name, producer, C, rows):
                s.y, s.cout, s.rows = producer.y, C, rows
                s.ps, s.pq = eng.part_sum, eng.part_sq
                s.fsum = producer.fsum
                _bn_views(s, eng, bn_name, C)
            stat_rows = 0

        F.append(lambda training: stem.fprop(self, True))
        x, H, W, producer = stem.y, 32, 32, stem
        nblk = len(spec["blocks"])
        if spec["blocks"] and "planes" in spec["blocks"][0]:
            return self._build_preact_bottleneck(stem, PreBN)
        for bi, b in enumerate(spec["blocks"]):
            name, cin, cout, s = b["name"], b["cin"], b["cout"], b["stride"]
            bn1 = PreBN(self, name + ".bn1", producer, cin, N * H * W)
            a = self.new_act(N, H, W, cin)
            c1 = ConvUnit(self, name + ".conv1", name + ".bn2", a, cin, cout, 3, s, H, W)
            Ho, Wo = H // s, W // s
            a2 = self.new_act(N, Ho, Wo, cout)
            c2 = ConvUnit(self, name + ".conv2", None, a2, cout, cout, 3, 1, Ho, Wo)
            cs = ConvUnit(self, name + ".shortcut.0", None, a, cin, cout, 1, s, H, W) if b["shortcut"] else None
            last = bi == nblk - 1
            self.units += [c1, c2] + ([cs] if cs else [])
            blocks.append((bn1, a, c1, a2, c2, cs, x))

            def fwd(training, bn1=bn1, a=a, c1=c1, a2=a2, c2=c2, cs=cs, x=x, producer=producer, last=last):
                bn1.stat_rows = producer.stat_rows
                bn1.bn_apply(self, training, a)
                if cs is not None:          # the shortcut conv overlaps conv1 on the side stream
                    self.run_side(lambda: cs.fprop(self, stats=False))
                c1.fprop(self)
                c1.bn_apply(self, training, a2)
                if cs is not None:
                    self.join_side()
                c2.fprop(self, stats=not last, add_src=(cs.y if cs is not None else x))
            F.append(fwd)
            x, H, W, producer = c2.y, Ho, Wo, c2
        fwd_head, bwd_head, dout = self._build_head(x, H, W, "linear.weight", "linear.bias")
        F.append(fwd_head)
        Bk = [bwd_head]
        for (bn1, a, c1, a2, c2, cs, xin) in reversed(blocks):
            da2 = self.new_act(*a2.shape)
            dy1 = self.new_act(*c1.y.shape)
            da = self.new_act(*a.shape)
            dx = self.new_act(*xin.shape)

            def bwd(bn1=bn1, a=a, c1=c1, a2=a2, c2=c2, cs=cs, dout=dout, da2=da2, dy1=dy1, da=da, dx=dx):
                c2.wgrad(dout)                              # d(conv2 out) = d(block out)
                if FUSE_DGRAD_BN:                           # a2 = relu(bn2(conv1 out))
                    c1.bn_backward_from_partial(self, da2, dy1, c2.dgrad_bn(self, dout, da2, c1))
                else:
                    c2.dgrad(dout, da2)
                    c1.bn_backward(self, da2, a2, da2, dy1, relu_from_y=True)
                c1.wgrad(dy1)
                if cs is not None:
                    cs.wgrad(dout)
                    c1.dgrad_with_shortcut(dy1, dout, cs, da)
                    bn1.bn_backward(self, da, a, da, dx, relu_from_y=True)    # a = relu(bn1(x)); no identity path
                else:
                    c1.dgrad(dy1, da)
                    bn1.bn_backward(self, da, a, da, dx, add=dout, relu_from_y=True)   # + gradient through the identity shortcut
            Bk.append(bwd)
            dout = dx
        Bk.append(lambda dout=dout: stem.wgrad(dout))
        self.bwd_ops = Bk

    def _build_preact_bottleneck(self, stem, PreBN):
        """models/preact_resnet.py:54-61 (PreActResNet-50/101/152): the pre-activation pattern of _build_preact with
        three convs; conv3's epilogue adds the shortcut and emits the statistics of the next block's bn1."""
        spec, N = self.spec, self.N
        F, blocks = self.fwd_ops, []
        x, H, W, producer = stem.y, 32, 32, stem
        nblk = len(spec["blocks"])
        for bi, b in enumerate(spec["blocks"]):
            name, cin, pl, cout, s = b["name"], b["cin"], b["planes"], b["cout"], b["stride"]
            bn1 = PreBN(self, name + ".bn1", producer, cin, N * H * W)
            a = self.new_act(N, H, W, cin)
            c1 = ConvUnit(self, name + ".conv1", name + ".bn2", a, cin, pl, 1, 1, H, W)
            a2 = self.new_act(N, H, W, pl)
            c2 = ConvUnit(self, name + ".conv2", name + ".bn3", a2, pl, pl, 3, s, H, W)
            Ho, Wo = H // s, W // s
            a3 = self.new_act(N, Ho, Wo, pl)
            c3 = ConvUnit(self, name + ".conv3", None, a3, pl, cout, 1, 1, Ho, Wo)
            cs = ConvUnit(self, name + ".shortcut.0", None, a, cin, cout, 1, s, H, W) if b["shortcut"] else None
            last = bi == nblk - 1
            self.units += [c1, c2, c3] + ([cs] if cs else [])
            blocks.append((bn1, a, c1, a2, c2, a3, c3, cs, x))

            def fwd(training, bn1=bn1, a=a, c1=c1, a2=a2, c2=c2, a3=a3, c3=c3, cs=cs, x=x, producer=producer, last=last):
                bn1.stat_rows = producer.stat_rows
                bn1.bn_apply(self, training, a)
                if cs is not None:
                    self.run_side(lambda: cs.fprop(self, stats=False))
                c1.fprop(self)
                c1.bn_apply(self, training, a2)
                c2.fprop(self)
                c2.bn_apply(self, training, a3)
                if cs is not None:
                    self.join_side()
                c3.fprop(self, stats=not last, add_src=(cs.y if cs is not None else x))
            F.append(fwd)
            x, H, W, producer = c3.y, Ho, Wo, c3
        fwd_head, bwd_head, dout = self._build_head(x, H, W, "linear.weight", "linear.bias")
        F.append(fwd_head)
        Bk = [bwd_head]
        for (bn1, a, c1, a2, c2, a3, c3, cs, xin) in reversed(blocks):
            da3, dy2, da2, dy1 = (self.new_act(*t.shape) for t in (a3, c2.y, a2, c1.y))
            da = self.new_act(*a.shape)
            dx = self.new_act(*xin.shape)

            def bwd(bn1=bn1, a=a, c1=c1, a2=a2, c2=c2, a3=a3, c3=c3, cs=cs, dout=dout, da3=da3, dy2=dy2, da2=da2,
                    dy1=dy1, da=da, dx=dx):
                c3.wgrad(dout)                              # d(conv3 out) = d(block out)
                if FUSE_DGRAD_BN:
                    c2.bn_backward_from_partial(self, da3, dy2, c3.dgrad_bn(self, dout, da3, c2))
                else:
                    c3.dgrad(dout, da3)
                    c2.bn_backward(self, da3, a3, da3, dy2, relu_from_y=True)
                c2.wgrad(dy2)
                if FUSE_DGRAD_BN:
                    c1.bn_backward_from_partial(self, da2, dy1, c2.dgrad_bn(self, dy2, da2, c1))
                else:
                    c2.dgrad(dy2, da2)
                    c1.bn_backward(self, da2, a2, da2, dy1, relu_from_y=True)
                c1.wgrad(dy1)
                if cs is not None:
                    cs.wgrad(dout)
                    cs.dgrad(dout, da)
                    c1.dgrad(dy1, da, add_src=da)
                    bn1.bn_backward(self, da, a, da, dx, relu_from_y=True)
                else:
                    c1.dgrad(dy1, da)
                    bn1.bn_backward(self, da, a, da, dx, add=dout, relu_from_y=True)   # + the identity shortcut
            Bk.append(bwd)
            dout = dx
        Bk.append(lambda dout=dout: stem.wgrad(dout))
        self.bwd_ops = Bk

    # ------------------------------------------------------------------ SENet plan (models/senet.py)
    def _build_senet(self):
        """models/senet.py:62-78,97-107: the pre-activation block of _build_preact with squeeze-and-excitation between
        conv2 and the residual add. Two things differ from PreActResNet: the stem has its own BN + ReLU, and a block's
        output (conv2 out * w + shortcut) is written by the SE scale pass, not by a conv epilogue — so the statistics
        the NEXT block's bn1 needs are taken from the materialised tensor (b2c_bn_bwd_reduce_acc with dz = y = the
        tensor accumulates [sum | sum of squares] into one row; b2c_bn_finalize reads it as a single partial row)."""
        spec, N = self.spec, self.N
        self.x0 = self.new_act(N, 32, 32, 64)
        stem = ConvUnit(self, "conv1", "bn1", self.x0, 3, 64, 3, 1, 32, 32, stem=True)
        a0 = self.new_act(N, 32, 32, 64)
        self.units = [stem]
        F, blocks = self.fwd_ops, []
        eng = self

        class Source:
            """A materialised tensor whose per-channel statistics feed a BatchNorm."""
            def __init__(s, y, C):
                s.y, s.C = y, C
                s.fsum = eng.new_stat(2 * C, forward=True)

            def take_stats(s):
                rows = s.y.numel() // s.C
                ops.bn_bwd_reduce_acc(s.y, None, s.y, None, rows, s.C, s.fsum)

        class PreBN(_BNOps):
            def __init__(s, bn_name, src, rows):
                C = src.C
                s.y, s.cout, s.rows = src.y, C, rows
                s.fsum = src.fsum
                s.ps, s.pq, s.stat_rows = src.fsum[:C], src.fsum[C:], 1
                _bn_views(s, eng, bn_name, C)

        def fwd_stem(training):
            stem.fprop(self, True)
            stem.bn_apply(self, training, a0)
        F.append(fwd_stem)
        src = Source(a0, 64)
        F.append(lambda training, src=src: src.take_stats() if training else None)
        x, H, W = a0, 32, 32
        max_c = max(b["cout"] for b in spec["blocks"])
        se_ds, se_dp = self.new_f32(N, max_c), self.new_f32(N, max_c)
        se_dz1 = self.new_f32(N, max_c // 16)
        nblk = len(spec["blocks"])
        for bi, b in enumerate(spec["blocks"]):
            name, cin, cout, s = b["name"], b["cin"], b["cout"], b["stride"]
            bn1 = PreBN(name + ".bn1", src, N * H * W)
            a = self.new_act(N, H, W, cin)
            c1 = ConvUnit(self, name + ".conv1", name + ".bn2", a, cin, cout, 3, s, H, W)
            Ho, Wo = H // s, W // s
            a2 = self.new_act(N, Ho, Wo, cout)
            c2 = ConvUnit(self, name + ".conv2", None, a2, cout, cout, 3, 1, Ho, Wo)
            cs = ConvUnit(self, name + ".shortcut.0", None, a, cin, cout, 1, s, H, W) if b["shortcut"] else None
            R = cout // 16
            se = dict(R=R, hw=Ho * Wo, C=cout,
                      w1=self.param_view(name + ".fc1.weight"), b1=self.param_view(name + ".fc1.bias"),
                      w2=self.param_view(name + ".fc2.weight"), b2=self.param_view(name + ".fc2.bias"),
                      dw1=self.grad_view(name + ".fc1.weight"), db1=self.grad_view(name + ".fc1.bias"),
                      dw2=self.grad_view(name + ".fc2.weight"), db2=self.grad_view(name + ".fc2.bias"),
                      pooled=self.new_f32(N, cout), hidden=self.new_f32(N, R), scale=self.new_f32(N, cout))
            out = self.new_act(N, Ho, Wo, cout)
            last = bi == nblk - 1
            nsrc = None if last else Source(out, cout)
            self.units += [c1, c2] + ([cs] if cs else [])
            blocks.append((bn1, a, c1, a2, c2, cs, se, x))

            def fwd(training, bn1=bn1, a=a, c1=c1, a2=a2, c2=c2, cs=cs, se=se, out=out, x=x, nsrc=nsrc):
                bn1.bn_apply(self, training, a)
                if cs is not None:          # the shortcut conv overlaps conv1 on the side stream
                    self.run_side(lambda: cs.fprop(self, stats=False))
                c1.fprop(self)
                c1.bn_apply(self, training, a2)
                c2.fprop(self, stats=False)
                if cs is not None:
                    self.join_side()
                ops.se_fwd(c2.y, N, se["hw"], se["C"], se["C"], se["R"], se["w1"], se["b1"], se["w2"], se["b2"],
                           se["pooled"], se["hidden"], se["scale"], out, add=(cs.y if cs is not None else x))
                if nsrc is not None and training:
                    nsrc.take_stats()
            F.append(fwd)
            x, H, W, src = out, Ho, Wo, nsrc
        fwd_head, bwd_head, dout = self._build_head(x, H, W, "linear.weight", "linear.bias")
        F.append(fwd_head)
        Bk = [bwd_head]
        for (bn1, a, c1, a2, c2, cs, se, xin) in reversed(blocks):
            dy2 = self.new_act(*c2.y.shape)
            da2 = self.new_act(*a2.shape)
            dy1 = self.new_act(*c1.y.shape)
            da = self.new_act(*a.shape)
            dx = self.new_act(*xin.shape)

            def bwd(bn1=bn1, a=a, c1=c1, a2=a2, c2=c2, cs=cs, se=se, dout=dout, dy2=dy2, da2=da2, dy1=dy1, da=da,
                    dx=dx):
                C, R = se["C"], se["R"]
                ops.se_bwd(dout, c2.y, N, se["hw"], C, C, R, se["w1"], se["w2"], se["pooled"], se["hidden"],
                           se["scale"], se_ds.view(-1)[:N * C], se_dz1.view(-1)[:N * R], se_dp.view(-1)[:N * C],
                           se["dw1"], se["db1"], se["dw2"], se["db2"], dy2)
                c2.wgrad(dy2)
                if FUSE_DGRAD_BN:                           # a2 = relu(bn2(conv1 out))
                    c1.bn_backward_from_partial(self, da2, dy1, c2.dgrad_bn(self, dy2, da2, c1))
                else:
                    c2.dgrad(dy2, da2)
                    c1.bn_backward(self, da2, a2, da2, dy1, relu_from_y=True)
                c1.wgrad(dy1)
                if cs is not None:
                    cs.wgrad(dout)
                    c1.dgrad_with_shortcut(dy1, dout, cs, da)
                    bn1.bn_backward(self, da, a, da, dx, relu_from_y=True)
                else:
                    c1.dgrad(dy1, da)
                    bn1.bn_backward(self, da, a, da, dx, add=dout, relu_from_y=True)   # + the identity shortcut
            Bk.append(bwd)
            dout = dx
        dy0 = self.new_act(*stem.y.shape)

        def bwd_stem(dout=dout):
            stem.bn_backward(self, dout, a0, dout, dy0, relu_from_y=True)
            stem.wgrad(dy0)
        Bk.append(bwd_stem)
        self.bwd_ops = Bk

    # ------------------------------------------------------------------ GoogLeNet plan (models/googlenet.py)
    def _build_googlenet(self):
        """A module's output is the LIST of its four branch outputs; the next module's 1x1 convs read the list through
        one tensor map per tensor (CatConvUnit), its pool branch max-pools each tensor on its own, and each tensor's
        gradient is the sum of what the four consumers send back (first writer plain, the others accumulate).
        MaxPool(3, 2, 1) between stages pools every tensor of the list; only the head, which averages the whole 8x8 map
        into Linear(1024, 10), sees a materialised concatenation (b2c_copy_channels)."""
        spec, N = self.spec, self.N
        self.x0 = self.new_act(N, 32, 32, 64)
        stem = ConvUnit(self, "pre_layers.0", "pre_layers.1", self.x0, 3, 192, 3, 1, 32, 32, stem=True)
        a0 = self.new_act(N, 32, 32, 192)
        self.units = [stem]
        F = self.fwd_ops
        eng = self

        def u8(t):
            return torch.zeros(t.shape, dtype=torch.uint8, device=self.device)

        class Inception:
            def __init__(I, d, srcs, H, W):
                name = d["name"]
                I.H, I.W = H, W
                I.xs, I.cs = [t for t, _ in srcs], [c for _, c in srcs]
                act = lambda c: eng.new_act(N, H, W, _pad64(c))
                I.b1 = CatConvUnit(eng, name + ".b1.0", name + ".b1.1", I.xs, I.cs, d["n1"], H, W)
                I.y1 = act(d["n1"])
                I.b2r = CatConvUnit(eng, name + ".b2.0", name + ".b2.1", I.xs, I.cs, d["n3r"], H, W)
                I.a2r = act(d["n3r"])
                I.b2 = ConvUnit(eng, name + ".b2.3", name + ".b2.4", I.a2r, d["n3r"], d["n3"], 3, 1, H, W)
                I.y2 = act(d["n3"])
                I.b3r = CatConvUnit(eng, name + ".b3.0", name + ".b3.1", I.xs, I.cs, d["n5r"], H, W)
                I.a3r = act(d["n5r"])
                I.b3a = ConvUnit(eng, name + ".b3.3", name + ".b3.4", I.a3r, d["n5r"], d["n5"], 3, 1, H, W)
                I.a3a = act(d["n5"])
                I.b3b = ConvUnit(eng, name + ".b3.6", name + ".b3.7", I.a3a, d["n5"], d["n5"], 3, 1, H, W)
                I.y3 = act(d["n5"])
                I.mp = [(eng.new_act(*x.shape), u8(x)) for x in I.xs]
                I.b4 = CatConvUnit(eng, name + ".b4.1", name + ".b4.2", [m for m, _ in I.mp], I.cs, d["npool"], H, W)
                I.y4 = act(d["npool"])
                I.out = [(I.y1, d["n1"]), (I.y2, d["n3"]), (I.y3, d["n5"]), (I.y4, d["npool"])]
                eng.units += [I.b1, I.b2r, I.b2, I.b3r, I.b3a, I.b3b, I.b4]

            def fwd(I, training):
                for unit, out in ((I.b1, I.y1), (I.b2r, I.a2r), (I.b2, I.y2), (I.b3r, I.a3r), (I.b3a, I.a3a),
                                  (I.b3b, I.y3)):
                    unit.fprop(eng)
                    unit.bn_apply(eng, training, out)
                for x, (m, arg) in zip(I.xs, I.mp):
                    ops.maxpool3x3_fwd(x, m, arg, 1)
                I.b4.fprop(eng)
                I.b4.bn_apply(eng, training, I.y4)

            def plan_bwd(I, gouts, ops_list, after=None):
                """gouts: gradient buffers of y1..y4. Appends the module's backward; returns the gradient buffers of
                its input tensors. after: extra work for the same backward op (the stage max-pool's backward)."""
                g1, g2, g3, g4 = gouts
                new = lambda t: eng.new_act(*t.shape)
                dsrc = [new(x) for x in I.xs]
                dmp = [new(x) for x in I.xs]
                dy1, dy2r, dy2, dy3r, dy3a, dy3b, dy4 = (new(u.y) for u in (I.b1, I.b2r, I.b2, I.b3r, I.b3a, I.b3b, I.b4))
                da2r, da3r, da3a = new(I.a2r), new(I.a3r), new(I.a3a)
                n = len(I.xs)

                def chain(conv, bn_unit, dy, da, dy_prev):
                    """dy -> through `conv` -> relu(bn(bn_unit.y)) backward -> dy_prev."""
                    if FUSE_DGRAD_BN:
                        bn_unit.bn_backward_from_partial(eng, da, dy_prev, conv.dgrad_bn(eng, dy, da, bn_unit))
                    else:
                        conv.dgrad(dy, da)
                        bn_unit.bn_backward(eng, da, None, da, dy_prev, relu_from_y=True)

                def run():
                    I.b1.bn_backward(eng, g1, I.y1, g1, dy1, relu_from_y=True)
                    I.b1.wgrad(dy1)
                    for t in range(n):
                        I.b1.dgrad(dy1, t, dsrc[t])
                    I.b2.bn_backward(eng, g2, I.y2, g2, dy2, relu_from_y=True)
                    I.b2.wgrad(dy2)
                    chain(I.b2, I.b2r, dy2, da2r, dy2r)
                    I.b2r.wgrad(dy2r)
                    for t in range(n):
                        I.b2r.dgrad(dy2r, t, dsrc[t], add_src=dsrc[t])
                    I.b3b.bn_backward(eng, g3, I.y3, g3, dy3b, relu_from_y=True)
                    I.b3b.wgrad(dy3b)
                    chain(I.b3b, I.b3a, dy3b, da3a, dy3a)
                    I.b3a.wgrad(dy3a)
                    chain(I.b3a, I.b3r, dy3a, da3r, dy3r)
                    I.b3r.wgrad(dy3r)
                    for t in range(n):
                        I.b3r.dgrad(dy3r, t, dsrc[t], add_src=dsrc[t])
                    I.b4.bn_backward(eng, g4, I.y4, g4, dy4, relu_from_y=True)
                    I.b4.wgrad(dy4)
                    for t in range(n):
                        I.b4.dgrad(dy4, t, dmp[t])
                        ops.maxpool3x3_bwd(dmp[t], I.mp[t][1], dsrc[t], 1, add=dsrc[t])
                    if after is not None:
                        after(dsrc)
                ops_list.append(run)
                return dsrc

        def fwd_stem(training):
            stem.fprop(self, True)
            stem.bn_apply(self, training, a0)
        F.append(fwd_stem)
        srcs, H, W = [(a0, 192)], 32, 32
        plan = []          # ("inc", Inception) | ("pool", [(x, pooled, arg)], H, W)
        for d in spec["modules"]:
            if d == "maxpool":
                pooled = [(x, self.new_act(N, H // 2, W // 2, x.shape[-1]),
                           torch.zeros(N, H // 2, W // 2, x.shape[-1], dtype=torch.uint8, device=self.device), c)
                          for x, c in srcs]
                plan.append(("pool", pooled))
                F.append(lambda training, pooled=pooled: [ops.maxpool3x3_fwd(x, y, arg, 2) for x, y, arg, _ in pooled])
                srcs, H, W = [(y, c) for _, y, _, c in pooled], H // 2, W // 2
            else:
                inc = Inception(d, srcs, H, W)
                plan.append(("inc", inc))
                F.append(inc.fwd)
                srcs = inc.out
        # head: the one place the concatenation exists
        ctot = sum(c for _, c in srcs)
        cat = self.new_act(N, H, W, ctot)
        final = srcs

        def fwd_cat(training):
            off = 0
            for t, c in final:
                ops.copy_channels(t, 0, cat, off, c)
                off += c
        F.append(fwd_cat)
        fwd_head, bwd_head, dcat = self._build_head(cat, H, W, "linear.weight", "linear.bias")
        F.append(fwd_head)
        gouts = [self.new_act(*t.shape) for t, _ in final]

        def bwd_head_split(gouts=gouts):
            bwd_head()
            off = 0
            for g, (t, c) in zip(gouts, final):
                ops.copy_channels(dcat, off, g, 0, c)
                off += c
        Bk = [bwd_head_split]
        for i in range(len(plan) - 1, -1, -1):
            kind, item = plan[i]
            if kind == "pool":
                continue
            after, gprev = None, None
            if i > 0 and plan[i - 1][0] == "pool":      # this module reads max-pooled tensors: un-pool its input gradients
                pooled = plan[i - 1][1]
                gprev = [self.new_act(*x.shape) for x, _, _, _ in pooled]

                def after(dsrc, pooled=pooled, gprev=gprev):
                    for (x, y, arg, c), dpool, gx in zip(pooled, dsrc, gprev):
                        ops.maxpool3x3_bwd(dpool, arg, gx, 2)
            sub = []
            dsrc = item.plan_bwd(gouts, sub, after)
            Bk.append(lambda sub=sub: [f() for f in sub])
            gouts = gprev if gprev is not None else dsrc
        dy0 = self.new_act(*stem.y.shape)
        d0 = gouts[0]

        def bwd_stem():
            stem.bn_backward(self, d0, a0, d0, dy0, relu_from_y=True)
            stem.wgrad(dy0)
        Bk.append(bwd_stem)
        self.bwd_ops = Bk

    # ------------------------------------------------------------------ Bottleneck ResNet plan
    def _build_bottleneck(self):
        """models/resnet.py:87-92: relu(bn1(conv1 1x1)) -> relu(bn2(conv2 3x3, stride)) -> bn3(conv3 1x1, x4)
        -> += shortcut(x) (identity, or 1x1 conv(stride) + BN) -> relu. ResNet-50/101/152 (main_dist.py:136)."""
        spec, N = self.spec, self.N
        self.x0 = self.new_act(N, 32, 32, 64)
        self.stem_taps = spec.get("stem_ksize", 3) ** 2       # which operand the input kernels build (trainer, accelerate)
        stem = ConvUnit(self, "conv1", "bn1", self.x0, 3, 64, spec.get("stem_ksize", 3), 1, 32, 32, stem=True)
        a0 = self.new_act(N, 32, 32, 64)
        self.units = [stem]
        F, blocks = self.fwd_ops, []

        def fwd_stem(training):
            stem.fprop(self, True)
            stem.bn_apply(self, training, a0)
        F.append(fwd_stem)
        x, H, W = a0, 32, 32
        for b in spec["blocks"]:
            name, cin, pl, cout, s = b["name"], b["cin"], b["planes"], b["cout"], b["stride"]
            c1 = ConvUnit(self, name + ".conv1", name + ".bn1", x, cin, pl, 1, 1, H, W)
            a1 = self.new_act(N, H, W, _pad64(pl))          # (RegNetX widths are off the 64-channel layout grid)
            c2 = ConvUnit(self, name + ".conv2", name + ".bn2", a1, pl, pl, 3, s, H, W)
            Ho, Wo = H // s, W // s
            a2 = self.new_act(N, Ho, Wo, _pad64(pl))
            se = None
            if b.get("se", 0):          # RegNetY: squeeze-and-excitation between relu(bn2(.)) and conv3 (regnet.py:60-61)
                sn, ld, R = name + ".se", _pad64(pl), b["se"]
                se = dict(R=R, hw=Ho * Wo, C=pl, ld=ld, out=self.new_act(N, Ho, Wo, ld),
                          w1=self.param_view(sn + ".se1.weight"), b1=self.param_view(sn + ".se1.bias"),
                          w2=self.param_view(sn + ".se2.weight"), b2=self.param_view(sn + ".se2.bias"),
                          dw1=self.grad_view(sn + ".se1.weight"), db1=self.grad_view(sn + ".se1.bias"),
                          dw2=self.grad_view(sn + ".se2.weight"), db2=self.grad_view(sn + ".se2.bias"),
                          pooled=self.new_f32(N, ld), hidden=self.new_f32(N, R), scale=self.new_f32(N, ld),
                          ds=self.new_f32(N * ld), dp=self.new_f32(N * ld), dz1=self.new_f32(N * R))
            c3 = ConvUnit(self, name + ".conv3", name + ".bn3", se["out"] if se else a2, pl, cout, 1, 1, Ho, Wo)
            cs = ConvUnit(self, name + ".shortcut.0", name + ".shortcut.1", x, cin, cout, 1, s, H, W,
                          private_stats=True) if b["shortcut"] else None
            out = self.new_act(N, Ho, Wo, _pad64(cout))
            bits = torch.zeros(out.numel() // 8, dtype=torch.uint8, device=self.device) if RELU_BITS else None
            self.units += [c1, c2, c3] + ([cs] if cs else [])
            blocks.append((c1, a1, c2, a2, c3, cs, out, x, bits, se))

            def fwd(training, c1=c1, a1=a1, c2=c2, a2=a2, c3=c3, cs=cs, out=out, x=x, bits=bits, se=se):
                if cs is not None:          # shortcut branch on the side stream (own statistics scratch)
                    self.run_side(lambda: (cs.fprop(self), None if (FUSED_FINALIZE and training) else cs.bn_stats(self, training)))
                c1.fprop(self)
                c1.bn_apply(self, training, a1)
                c2.fprop(self)
                c2.bn_apply(self, training, a2)
                if se is not None:
                    ops.se_fwd(a2, N, se["hw"], se["C"], se["ld"], se["R"], se["w1"], se["b1"], se["w2"], se["b2"],
                               se["pooled"], se["hidden"], se["scale"], se["out"])
                c3.fprop(self)
                if cs is not None:
                    self.join_side()
                    c3.bn_apply(self, training, out, res_unit=cs, relu_bits=bits)
                else:
                    c3.bn_apply(self, training, out, res=x, relu_bits=bits)
            F.append(fwd)
            x, H, W = out, Ho, Wo
        fwd_head, bwd_head, dout = self._build_head(x, H, W, "linear.weight", "linear.bias")
        F.append(fwd_head)
        Bk = [bwd_head]
        for (c1, a1, c2, a2, c3, cs, out, xin, bits, se) in reversed(blocks):
            dy3 = self.new_act(*c3.y.shape)
            da2s = self.new_act(*a2.shape) if se is not None else None
            dys = self.new_act(*cs.y.shape) if cs is not None else None
            da2 = self.new_act(*a2.shape)
            dy2 = self.new_act(*c2.y.shape)
            da1 = self.new_act(*a1.shape)
            dy1 = self.new_act(*c1.y.shape)
            dx = self.new_act(*xin.shape)

            def bwd(c1=c1, a1=a1, c2=c2, a2=a2, c3=c3, cs=cs, out=out, dout=dout, dy3=dy3, dys=dys, da2=da2,
                    dy2=dy2, da1=da1, dy1=dy1, dx=dx, bits=bits, se=se, da2s=da2s):
                c3.bn_backward(self, dout, out, dout, dy3, relu_bits=bits)      # mask in place -> dz
                if cs is not None:
                    cs.bn_backward(self, dout, None, None, dys)
                c3.wgrad(dy3)
                if se is not None:          # conv3 read a2 * w: back through the excitation, then the plain BN backward
                    c3.dgrad(dy3, da2s)
                    ops.se_bwd(da2s, a2, N, se["hw"], se["C"], se["ld"], se["R"], se["w1"], se["w2"], se["pooled"],
                               se["hidden"], se["scale"], se["ds"], se["dz1"], se["dp"], se["dw1"], se["db1"],
                               se["dw2"], se["db2"], da2)
                    c2.bn_backward(self, da2, a2, da2, dy2, relu_from_y=True)
                elif FUSE_DGRAD_BN:
                    c2.bn_backward_from_partial(self, da2, dy2, c3.dgrad_bn(self, dy3, da2, c2))
                else:
                    c3.dgrad(dy3, da2)
                    c2.bn_backward(self, da2, a2, da2, dy2, relu_from_y=True)
                c2.wgrad(dy2)
                if FUSE_DGRAD_BN and not c2.sliced:
                    c1.bn_backward_from_partial(self, da1, dy1, c2.dgrad_bn(self, dy2, da1, c1))
                else:       # (a conv run group by group has no fused dgrad + BatchNorm-backward epilogue)
                    c2.dgrad(dy2, da1)
                    c1.bn_backward(self, da1, a1, da1, dy1, relu_from_y=True)
                c1.wgrad(dy1)
                if cs is not None:
                    cs.wgrad(dys)
                    cs.dgrad(dys, dx)
                    c1.dgrad(dy1, dx, add_src=dx)
                else:
                    c1.dgrad(dy1, dx, add_src=dout)
            Bk.append(bwd)
            dout = dx
        dy0 = self.new_act(*stem.y.shape)

        def bwd_stem(dout=dout):
            stem.bn_backward(self, dout, a0, dout, dy0, relu_from_y=True)
            stem.wgrad(dy0)
        Bk.append(bwd_stem)
        self.bwd_ops = Bk

    # ------------------------------------------------------------------ MobileNetV2 plan
    def _build_mobilenetv2(self):
        """models/mobilenetv2.py:32-37,69-77: per block 1x1 expand + BN + ReLU -> depthwise 3x3 + BN + ReLU ->
        1x1 project + BN (linear); `out + shortcut(x)` only when stride == 1 (identity, or 1x1 conv + BN when the
        channel count changes). Head: 1x1 320->1280 + BN + ReLU, avg_pool2d(4), Linear(1280, 10)."""
        spec, N = self.spec, self.N
        self.x0 = self.new_act(N, 32, 32, 64)
        stem = ConvUnit(self, "conv1", "bn1", self.x0, 3, 32, 3, 1, 32, 32, stem=True)
        a0 = self.new_act(N, 32, 32, stem.cout)
        self.units = [stem]
        F, blocks = self.fwd_ops, []

        def fwd_stem(training):
            stem.fprop(self, True)
            stem.bn_apply(self, training, a0)
        F.append(fwd_stem)
        x, H, W = a0, 32, 32
        for b in spec["blocks"]:
            name, cin, cout, planes, s = b["name"], b["cin"], b["cout"], b["planes"], b["stride"]
            c1 = ConvUnit(self, name + ".conv1", name + ".bn1", x, cin, planes, 1, 1, H, W)
            fused_in = DW_FUSED_INPUT and ops.dwconv_input_transform_ok(H, W, c1.cout, s)
            a1 = None if fused_in else self.new_act(N, H, W, c1.cout)
            dw = DwUnit(self, name + ".conv2", name + ".bn2", a1, planes, s, H, W, src_unit=c1 if fused_in else None)
            Ho, Wo = H // s, W // s
            a2 = self.new_act(N, Ho, Wo, dw.cout)
            c3 = ConvUnit(self, name + ".conv3", name + ".bn3", a2, planes, cout, 1, 1, Ho, Wo)
            cs = ConvUnit(self, name + ".shortcut.0", name + ".shortcut.1", x, cin, cout, 1, 1, H, W,
                          private_stats=True) if b["shortcut"] else None
            identity = s == 1 and cs is None
            out = self.new_act(N, Ho, Wo, c3.cout)
            self.units += [c1, dw, c3] + ([cs] if cs else [])
            blocks.append((c1, a1, dw, a2, c3, cs, identity, out, x))

            def fwd(training, c1=c1, a1=a1, dw=dw, a2=a2, c3=c3, cs=cs, identity=identity, out=out, x=x):
                if cs is not None:          # shortcut branch on the side stream (own statistics scratch)
                    self.run_side(lambda: (cs.fprop(self), None if (FUSED_FINALIZE and training) else cs.bn_stats(self, training)))
                c1.fprop(self)
                if a1 is None:          # relu(bn1(.)) is applied by the depthwise kernels while they load c1.y
                    c1.bn_scale_shift(self, training)
                else:
                    c1.bn_apply(self, training, a1)
                dw.fprop(self)
                dw.bn_apply(self, training, a2)
                c3.fprop(self)
                if cs is not None:
                    self.join_side()
                    c3.bn_apply(self, training, out, relu=False, res_unit=cs)
                else:
                    c3.bn_apply(self, training, out, relu=False, res=(x if identity else None))
            F.append(fwd)
            x, H, W = out, Ho, Wo
        ch = ConvUnit(self, "conv2", "bn2", x, 320, 1280, 1, 1, H, W)
        ah = self.new_act(N, H, W, ch.cout)
        self.units.append(ch)

        def fwd_last(training):
            ch.fprop(self)
            ch.bn_apply(self, training, ah)
        F.append(fwd_last)
        fwd_head, bwd_head, dout = self._build_head(ah, H, W, "linear.weight", "linear.bias")
        F.append(fwd_head)
        Bk = [bwd_head]
        dyh = self.new_act(*ch.y.shape)
        dxh = self.new_act(*x.shape)

        def bwd_last(dout=dout):
            ch.bn_backward(self, dout, ah, dout, dyh, relu_from_y=True)
            ch.wgrad(dyh)
            ch.dgrad(dyh, dxh)
        Bk.append(bwd_last)
        dout = dxh
        for (c1, a1, dw, a2, c3, cs, identity, out, xin) in reversed(blocks):
            dy3 = self.new_act(*c3.y.shape)
            dys = self.new_act(*cs.y.shape) if cs is not None else None
            da2 = self.new_act(*a2.shape)
            dy2 = self.new_act(*dw.y.shape)
            da1 = self.new_act(*c1.y.shape)
            dy1 = self.new_act(*c1.y.shape)
            dx = self.new_act(*xin.shape)

            def bwd(c1=c1, a1=a1, dw=dw, a2=a2, c3=c3, cs=cs, identity=identity, dout=dout, dy3=dy3, dys=dys,
                    da2=da2, dy2=dy2, da1=da1, dy1=dy1, dx=dx):
                c3.bn_backward(self, dout, None, None, dy3)          # block output has no ReLU
                if cs is not None:
                    cs.bn_backward(self, dout, None, None, dys)
                c3.wgrad(dy3)
                if FUSE_DGRAD_BN:
                    dw.bn_backward_from_partial(self, da2, dy2, c3.dgrad_bn(self, dy3, da2, dw))
                else:
                    c3.dgrad(dy3, da2)
                    dw.bn_backward(self, da2, a2, da2, dy2, relu_from_y=True)
                dw.wgrad(dy2)
                dw.dgrad(dy2, da1)
                c1.bn_backward(self, da1, a1, da1, dy1, relu_from_y=True)
                c1.wgrad(dy1)
                if cs is not None:
                    cs.wgrad(dys)
                    cs.dgrad(dys, dx)
                    c1.dgrad(dy1, dx, add_src=dx)
                else:
                    c1.dgrad(dy1, dx, add_src=(dout if identity else None))
            Bk.append(bwd)
            dout = dx
        dy0 = self.new_act(*stem.y.shape)

        def bwd_stem(dout=dout):
            stem.bn_backward(self, dout, a0, dout, dy0, relu_from_y=True)
            stem.wgrad(dy0)
        Bk.append(bwd_stem)
        self.bwd_ops = Bk

    # ------------------------------------------------------------------ MobileNet (v1) plan
    def _build_mobilenet(self):
        """models/mobilenet.py:20-23,47-52: per block relu(bn1(depthwise 3x3, stride)) -> relu(bn2(pointwise 1x1));
        stem conv3x3 + BN + ReLU; avg_pool2d(2) over the final 2x2 map; Linear(1024, 10)."""
        spec, N = self.spec, self.N
        self.x0 = self.new_act(N, 32, 32, 64)
        stem = ConvUnit(self, "conv1", "bn1", self.x0, 3, 32, 3, 1, 32, 32, stem=True)
        a0 = self.new_act(N, 32, 32, stem.cout)
        self.units = [stem]
        F, blocks = self.fwd_ops, []
        F.append(lambda training: (stem.fprop(self, True), stem.bn_apply(self, training, a0)))
        x, H, W = a0, 32, 32
        for b in spec["blocks"]:
            name, cin, cout, s = b["name"], b["cin"], b["cout"], b["stride"]
            dw = DwUnit(self, name + ".conv1", name + ".bn1", x, cin, s, H, W)
            Ho, Wo = H // s, W // s
            a1 = self.new_act(N, Ho, Wo, dw.cout)
            c2 = ConvUnit(self, name + ".conv2", name + ".bn2", a1, cin, cout, 1, 1, Ho, Wo)
            a2 = self.new_act(N, Ho, Wo, c2.cout)
            self.units += [dw, c2]
            blocks.append((dw, a1, c2, a2, x))

            def fwd(training, dw=dw, a1=a1, c2=c2, a2=a2):
                dw.fprop(self)
                dw.bn_apply(self, training, a1)
                c2.fprop(self)
                c2.bn_apply(self, training, a2)
            F.append(fwd)
            x, H, W = a2, Ho, Wo
        fwd_head, bwd_head, dout = self._build_head(x, H, W, "linear.weight", "linear.bias")
        F.append(fwd_head)
        Bk = [bwd_head]
        for (dw, a1, c2, a2, xin) in reversed(blocks):
            dy2 = self.new_act(*c2.y.shape)
            da1 = self.new_act(*a1.shape)
            dy1 = self.new_act(*dw.y.shape)
            dx = self.new_act(*xin.shape)

            def bwd(dw=dw, a1=a1, c2=c2, a2=a2, dout=dout, dy2=dy2, da1=da1, dy1=dy1, dx=dx):
                c2.bn_backward(self, dout, a2, dout, dy2, relu_from_y=True)
                c2.wgrad(dy2)
                if FUSE_DGRAD_BN:       # a1 = relu(bn1(depthwise out)): mask + sums from the pointwise dgrad's epilogue
                    dw.bn_backward_from_partial(self, da1, dy1, c2.dgrad_bn(self, dy2, da1, dw))
                else:
                    c2.dgrad(dy2, da1)
                    dw.bn_backward(self, da1, a1, da1, dy1, relu_from_y=True)
                dw.wgrad(dy1)
                dw.dgrad(dy1, dx)
            Bk.append(bwd)
            dout = dx
        dy0 = self.new_act(*stem.y.shape)

        def bwd_stem(dout=dout):
            stem.bn_backward(self, dout, a0, dout, dy0, relu_from_y=True)
            stem.wgrad(dy0)
        Bk.append(bwd_stem)
        self.bwd_ops = Bk

    # ------------------------------------------------------------------ execution
    def forward(self, training=True):
        if FUSED_FINALIZE or self.spec["family"] == "senet" or self.zero_stats_each_forward:
            self._stat_f.zero_()      # (senet: statistics of materialised tensors; grouped convs run group by group)
        for op in self.fwd_ops:
            op(training)
        return self.logits

    # ---- side stream for work off the critical path (weight gradients)
    def run_side(self, fn):
        """Run fn() on the side stream, ordered after everything enqueued on the current stream so far. Its results
        may be used only after join_side() / side_marker(). Work submitted this way is serialised among itself (one
        side stream), which is what the shared halo-wgrad scratch buffer needs."""
        if not ASYNC_WGRAD:
            fn()
            return
        if self._side is None:
            self._side = torch.cuda.Stream(device=self.device)
        main = torch.cuda.current_stream()
        ev = torch.cuda.Event()
        ev.record(main)
        self._side.wait_event(ev)
        with torch.cuda.stream(self._side):
            fn()
        self._side_dirty = True

    def side_marker(self):
        """Event after the side-stream work submitted so far (None if there is none outstanding)."""
        if not self._side_dirty:
            return None
        ev = torch.cuda.Event()
        ev.record(self._side)
        return ev

    def join_side(self):
        ev = self.side_marker()
        if ev is not None:
            torch.cuda.current_stream().wait_event(ev)
            self._side_dirty = False

    def backward(self):
        """Expects self.dlogits filled and self.G zeroed (conv wgrad accumulates)."""
        for op in self.bwd_ops:
            op()
        self.join_side()

    def backward_allreduce(self, buckets, ready_after, comm_stream, group=None, average=False):
        """backward() with the DDP Reducer's job interleaved (main_dist.py:141-144): as soon as the backward op
        that completes a bucket of the gradient arena has been enqueued, that bucket is all-reduced with NCCL on
        `comm_stream` while the remaining backward kernels keep running; returns after making the current stream
        wait for the last bucket. buckets / ready_after come from plan_grad_buckets(). average=False leaves the
        SUM in G (the trainer folds 1/world into its SGD kernel); True makes NCCL average, as DDP hands .grad
        to a torch optimizer."""
        import torch.distributed as dist
        main = torch.cuda.current_stream()
        op_kind = dist.ReduceOp.AVG if average else dist.ReduceOp.SUM
        for pos, op in enumerate(self.bwd_ops):
            op()
            bi = ready_after.get(pos)
            if bi is not None:
                a, b = buckets[bi]
                ev = torch.cuda.Event()
                ev.record(main)
                comm_stream.wait_event(ev)
                side = self.side_marker()              # weight gradients of this bucket run on the side stream
                if side is not None:
                    comm_stream.wait_event(side)
                with torch.cuda.stream(comm_stream):
                    dist.all_reduce(self.G[a:b], op=op_kind, group=group)
        self.join_side()
        done = torch.cuda.Event()
        done.record(comm_stream)
        main.wait_event(done)

    def zero_grad(self):
        self.G.zero_()
        if FUSED_FINALIZE:
            self._stat_b.zero_()

    def activation_bytes(self):
        return sum(t.numel() * t.element_size() for t in self._acts)

