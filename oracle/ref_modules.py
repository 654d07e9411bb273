"""nn.Module stand-ins with the reference's attribute names and construction order (TEST INFRASTRUCTURE ONLY).

The GPU box has no /root/reference, so tests of the drop-in surface (pytorch-cifar_b200/accelerate.py) cannot
import models/*.py there. These classes rebuild the same module trees — same attribute names, hence the same
state_dict keys and the same parameter-creation order, hence the same seeded initial weights — for
ResNet/BasicBlock (models/resnet.py:16-51,96-140), VGG (models/vgg.py:14-38) and PreActResNet/PreActBlock
(models/preact_resnet.py:12-34,64-94). tests/test_oracle.py pins them against the real reference modules
(identical keys, identical seeded weights, identical outputs). Class NAMES match the reference on purpose:
accelerate() recognises module trees structurally, by class and attribute names.
"""
import torch.nn as nn
import torch.nn.functional as F

from .nets import MOBILENETV2_CFG, VGG_CFG


def _conv(cin, cout, k, stride):
    return nn.Conv2d(cin, cout, kernel_size=k, stride=stride, padding=k // 2, bias=False)


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, cin, cout, stride):
        super().__init__()
        self.conv1, self.bn1 = _conv(cin, cout, 3, stride), nn.BatchNorm2d(cout)
        self.conv2, self.bn2 = _conv(cout, cout, 3, 1), nn.BatchNorm2d(cout)
        self.shortcut = nn.Sequential()
        if stride != 1 or cin != cout:
            self.shortcut = nn.Sequential(_conv(cin, cout, 1, stride), nn.BatchNorm2d(cout))

    def forward(self, x):
        y = F.relu(self.bn1(self.conv1(x)))
        return F.relu(self.bn2(self.conv2(y)) + self.shortcut(x))


class PreActBlock(nn.Module):
    expansion = 1

    def __init__(self, cin, cout, stride):
        super().__init__()
        self.bn1, self.conv1 = nn.BatchNorm2d(cin), _conv(cin, cout, 3, stride)
        self.bn2, self.conv2 = nn.BatchNorm2d(cout), _conv(cout, cout, 3, 1)
        if stride != 1 or cin != cout:
            self.shortcut = nn.Sequential(_conv(cin, cout, 1, stride))

    def forward(self, x):
        a = F.relu(self.bn1(x))
        skip = self.shortcut(a) if hasattr(self, "shortcut") else x
        return self.conv2(F.relu(self.bn2(self.conv1(a)))) + skip


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, cin, planes, stride):
        super().__init__()
        cout = 4 * planes
        self.conv1, self.bn1 = _conv(cin, planes, 1, 1), nn.BatchNorm2d(planes)
        self.conv2, self.bn2 = _conv(planes, planes, 3, stride), nn.BatchNorm2d(planes)
        self.conv3, self.bn3 = _conv(planes, cout, 1, 1), nn.BatchNorm2d(cout)
        self.shortcut = nn.Sequential()
        if stride != 1 or cin != cout:
            self.shortcut = nn.Sequential(_conv(cin, cout, 1, stride), nn.BatchNorm2d(cout))

    def forward(self, x):
        y = F.relu(self.bn2(self.conv2(F.relu(self.bn1(self.conv1(x))))))
        return F.relu(self.bn3(self.conv3(y)) + self.shortcut(x))


def _stages(block, counts):
    widths, cin, out = (64, 128, 256, 512), 64, []
    for i, (n, w) in enumerate(zip(counts, widths)):
        seq = []
        for j in range(n):
            seq.append(block(cin, w, (1 if i == 0 else 2) if j == 0 else 1))
            cin = w * block.expansion
        out.append(nn.Sequential(*seq))
    return out


class ResNet(nn.Module):
    def __init__(self, counts=(2, 2, 2, 2), num_classes=10, block=BasicBlock):
        super().__init__()
        self.conv1, self.bn1 = _conv(3, 64, 3, 1), nn.BatchNorm2d(64)
        self.layer1, self.layer2, self.layer3, self.layer4 = _stages(block, counts)
        self.linear = nn.Linear(512 * block.expansion, num_classes)

    def forward(self, x):
        y = F.relu(self.bn1(self.conv1(x)))
        y = self.layer4(self.layer3(self.layer2(self.layer1(y))))
        return self.linear(F.avg_pool2d(y, 4).flatten(1))


class PreActResNet(nn.Module):
    def __init__(self, counts=(2, 2, 2, 2), num_classes=10):
        super().__init__()
        self.conv1 = _conv(3, 64, 3, 1)
        self.layer1, self.layer2, self.layer3, self.layer4 = _stages(PreActBlock, counts)
        self.linear = nn.Linear(512, num_classes)

    def forward(self, x):
        y = self.layer4(self.layer3(self.layer2(self.layer1(self.conv1(x)))))
        return self.linear(F.avg_pool2d(y, 4).flatten(1))


class VGG(nn.Module):
    def __init__(self, name="VGG16"):
        super().__init__()
        seq, cin = [], 3
        for v in VGG_CFG[name]:
            if v == "M":
                seq.append(nn.MaxPool2d(kernel_size=2, stride=2))
            else:
                seq += [nn.Conv2d(cin, v, kernel_size=3, padding=1), nn.BatchNorm2d(v), nn.ReLU(inplace=True)]
                cin = v
        seq.append(nn.AvgPool2d(kernel_size=1, stride=1))
        self.features = nn.Sequential(*seq)
        self.classifier = nn.Linear(512, 10)

    def forward(self, x):
        return self.classifier(self.features(x).flatten(1))


class Block(nn.Module):
    """MobileNetV2 inverted residual (models/mobilenetv2.py:11-37)."""

    def __init__(self, cin, cout, expansion, stride):
        super().__init__()
        self.stride, p = stride, expansion * cin
        self.conv1, self.bn1 = _conv(cin, p, 1, 1), nn.BatchNorm2d(p)
        self.conv2 = nn.Conv2d(p, p, kernel_size=3, stride=stride, padding=1, groups=p, bias=False)
        self.bn2 = nn.BatchNorm2d(p)
        self.conv3, self.bn3 = _conv(p, cout, 1, 1), nn.BatchNorm2d(cout)
        self.shortcut = nn.Sequential()
        if stride == 1 and cin != cout:
            self.shortcut = nn.Sequential(_conv(cin, cout, 1, 1), nn.BatchNorm2d(cout))

    def forward(self, x):
        y = F.relu(self.bn2(self.conv2(F.relu(self.bn1(self.conv1(x))))))
        y = self.bn3(self.conv3(y))
        return y + self.shortcut(x) if self.stride == 1 else y


class MobileNetV2(nn.Module):
    def __init__(self, num_classes=10):
        super().__init__()
        self.conv1, self.bn1 = _conv(3, 32, 3, 1), nn.BatchNorm2d(32)
        seq, cin = [], 32
        for e, cout, n, stride in MOBILENETV2_CFG:
            for s in [stride] + [1] * (n - 1):
                seq.append(Block(cin, cout, e, s))
                cin = cout
        self.layers = nn.Sequential(*seq)
        self.conv2, self.bn2 = _conv(320, 1280, 1, 1), nn.BatchNorm2d(1280)
        self.linear = nn.Linear(1280, num_classes)

    def forward(self, x):
        y = self.layers(F.relu(self.bn1(self.conv1(x))))
        y = F.relu(self.bn2(self.conv2(y)))
        return self.linear(F.avg_pool2d(y, 4).flatten(1))


def ResNet18():
    return ResNet((2, 2, 2, 2))


def ResNet50():
    return ResNet((3, 4, 6, 3), block=Bottleneck)


def PreActResNet18():
    return PreActResNet((2, 2, 2, 2))
