"""Generates the golden fixtures in this directory FROM THE REFERENCE ITSELF.

Run in the build container only (it imports /root/reference/models and torchvision.transforms; neither
travels to the GPU box):   python tests/golden/make_golden.py
The fixtures pin oracle/ (tests/test_oracle.py) and are small enough to commit.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, "/root/reference")


def golden_resnet18():
    import models.resnet as ref_resnet           # the reference's own model code
    torch.manual_seed(0)
    net = ref_resnet.ResNet18()
    net.train()
    rs = np.random.RandomState(0)
    x = torch.from_numpy(rs.standard_normal((4, 3, 32, 32)).astype(np.float32))
    target = torch.from_numpy(rs.randint(0, 10, size=(4,)).astype(np.int64))
    logits = net(x)                              # main.py:102
    loss = torch.nn.CrossEntropyLoss()(logits, target)   # main.py:86,103
    loss.backward()                              # main.py:104
    out = dict(x=x.numpy(), target=target.numpy(), logits=logits.detach().numpy(), loss=np.float32(loss.item()))
    names, norms, heads, wsum = [], [], [], []
    for name, p in net.named_parameters():
        names.append(name)
        g = p.grad.detach().reshape(-1)
        norms.append(g.norm().item())
        heads.append(np.pad(g[:8].numpy(), (0, max(0, 8 - g.numel()))))
        wsum.append(p.detach().double().sum().item())
    out["param_names"] = np.array(names)
    out["grad_norms"] = np.array(norms, dtype=np.float64)
    out["grad_heads"] = np.stack(heads).astype(np.float32)
    out["weight_sums"] = np.array(wsum, dtype=np.float64)
    bn_names, bn_rm, bn_rv = [], [], []
    for name, b in net.named_buffers():
        if name.endswith("running_mean"):
            bn_names.append(name[:-len(".running_mean")])
            bn_rm.append(b.double().sum().item())
        if name.endswith("running_var"):
            bn_rv.append(b.double().sum().item())
    out["bn_names"] = np.array(bn_names)
    out["bn_rm_sums"] = np.array(bn_rm)
    out["bn_rv_sums"] = np.array(bn_rv)
    # eval-mode forward after the one training forward (running stats updated once)
    net.eval()
    with torch.no_grad():
        out["logits_eval"] = net(x).numpy()
    np.savez_compressed(os.path.join(HERE, "resnet18_seed0_n4.npz"), **out)
    print("resnet18: loss", loss.item(), "params", len(names))


def golden_other_nets():
    """VGG-16 and PreActResNet-18 (BASELINE configs 3, 4): init sums, logits, loss, gradient norms at N=4."""
    import models.vgg as ref_vgg
    import models.preact_resnet as ref_pre
    rs = np.random.RandomState(2)
    x = torch.from_numpy(rs.standard_normal((4, 3, 32, 32)).astype(np.float32))
    target = torch.from_numpy(rs.randint(0, 10, size=(4,)).astype(np.int64))
    out = dict(x=x.numpy(), target=target.numpy())
    import models.mobilenetv2 as ref_mb
    import models.dla_simple as ref_dla
    import models.mobilenet as ref_mb1
    for tag, ctor in (("vgg16", lambda: ref_vgg.VGG("VGG16")), ("preact18", ref_pre.PreActResNet18),
                      ("mobilenetv2", ref_mb.MobileNetV2), ("resnet50", __import__("models.resnet", fromlist=["x"]).ResNet50),
                      ("simpledla", ref_dla.SimpleDLA), ("mobilenet", ref_mb1.MobileNet)):
        torch.manual_seed(0)
        net = ctor()
        net.train()
        logits = net(x)
        loss = torch.nn.CrossEntropyLoss()(logits, target)
        loss.backward()
        names = [n for n, _ in net.named_parameters()]
        out[tag + "_names"] = np.array(names)
        out[tag + "_wsum"] = np.array([p.detach().double().sum().item() for _, p in net.named_parameters()])
        out[tag + "_gnorm"] = np.array([p.grad.norm().item() for _, p in net.named_parameters()])
        out[tag + "_logits"] = logits.detach().numpy()
        out[tag + "_loss"] = np.float32(loss.item())
        print(tag, "loss", loss.item(), "params", len(names))
    np.savez_compressed(os.path.join(HERE, "other_nets_seed0_n4.npz"), **out)


def golden_resnext():
    """ResNeXt29_2x64d / 32x4d (models/resnext.py:78-88) and RegNetX_200MF / 400MF (models/regnet.py:113-134), the
    networks with grouped convs, DLA (models/dla.py) and SENet18 (models/senet.py): the same quantities as golden_other_nets, in a file of
    their own."""
    import models.resnext as ref_rx
    import models.regnet as ref_rg
    rs = np.random.RandomState(2)
    x = torch.from_numpy(rs.standard_normal((4, 3, 32, 32)).astype(np.float32))
    target = torch.from_numpy(rs.randint(0, 10, size=(4,)).astype(np.int64))
    out = dict(x=x.numpy(), target=target.numpy())
    for tag, ctor in (("resnext29_2x64d", ref_rx.ResNeXt29_2x64d), ("resnext29_32x4d", ref_rx.ResNeXt29_32x4d),
                      ("regnetx_200mf", ref_rg.RegNetX_200MF), ("regnetx_400mf", ref_rg.RegNetX_400MF),
                      ("dla", __import__("models.dla", fromlist=["x"]).DLA),
                      ("senet18", __import__("models.senet", fromlist=["x"]).SENet18),
                      ("regnety_400mf", ref_rg.RegNetY_400MF),
                      ("googlenet", __import__("models.googlenet", fromlist=["x"]).GoogLeNet),
                      ("preact50", __import__("models.preact_resnet", fromlist=["x"]).PreActResNet50)):
        torch.manual_seed(0)
        net = ctor()
        net.train()
        logits = net(x)
        loss = torch.nn.CrossEntropyLoss()(logits, target)
        loss.backward()
        names = [n for n, _ in net.named_parameters()]
        out[tag + "_names"] = np.array(names)
        out[tag + "_wsum"] = np.array([p.detach().double().sum().item() for _, p in net.named_parameters()])
        out[tag + "_gnorm"] = np.array([p.grad.norm().item() for _, p in net.named_parameters()])
        out[tag + "_logits"] = logits.detach().numpy()
        out[tag + "_loss"] = np.float32(loss.item())
        print(tag, "loss", loss.item(), "params", len(names))
    np.savez_compressed(os.path.join(HERE, "grouped_nets_seed0_n4.npz"), **out)


def golden_augment():
    import torchvision.transforms as transforms
    from PIL import Image
    tf = transforms.Compose([                      # main.py:30-35 verbatim
        transforms.RandomCrop(32, padding=4),
        transforms.RandomHorizontalFlip(),
        transforms.ToTensor(),
        transforms.Normalize((0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010)),
    ])
    rs = np.random.RandomState(7)
    imgs = rs.randint(0, 256, size=(24, 32, 32, 3)).astype(np.uint8)
    torch.manual_seed(123)                          # num_workers=0 semantics: the global CPU generator
    outs = [tf(Image.fromarray(im)).numpy() for im in imgs]     # TV/datasets/cifar.py:118-121
    np.savez_compressed(os.path.join(HERE, "augment_seed123.npz"), images=imgs, seed=np.int64(123),
                        out=np.stack(outs).astype(np.float32))
    print("augment: 24 samples")


def golden_sgd():
    torch.manual_seed(5)
    p = torch.randn(257, requires_grad=True)
    p0 = p.detach().clone().numpy()
    opt = torch.optim.SGD([p], lr=0.1, momentum=0.9, weight_decay=5e-4)   # main.py:87-88
    sched = torch.optim.lr_scheduler.CosineAnnealingLR(opt, T_max=200)     # main.py:89
    grads, traj, lrs = [], [], []
    for step in range(4):
        g = torch.randn(257)
        grads.append(g.numpy().copy())
        p.grad = g.clone()
        opt.step()
        traj.append(p.detach().numpy().copy())
    for e in range(5):
        lrs.append(opt.param_groups[0]["lr"])
        sched.step()
    np.savez_compressed(os.path.join(HERE, "sgd_traj.npz"), p0=p0, grads=np.stack(grads), traj=np.stack(traj),
                        lrs=np.array(lrs, dtype=np.float64))
    print("sgd: 4 steps")


def golden_sampler():
    from torch.utils.data.distributed import DistributedSampler

    class _DS:
        def __len__(self):
            return 50000
    out = {}
    for world in (1, 2, 4, 8):
        for rank in (0, world - 1):
            s = DistributedSampler(_DS(), num_replicas=world, rank=rank)   # main_dist.py:110, epoch stays 0
            idx = np.array(list(iter(s)), dtype=np.int64)
            out[f"w{world}_r{rank}_head"] = idx[:64]
            out[f"w{world}_r{rank}_len"] = np.int64(len(idx))
            out[f"w{world}_r{rank}_sum"] = np.int64(idx.sum())
    np.savez_compressed(os.path.join(HERE, "sampler.npz"), **out)
    print("sampler ok")


if __name__ == "__main__":
    if sys.argv[1:] == ["resnext"]:          # only the fixture added last; the others stay byte-identical
        golden_resnext()
        sys.exit(0)
    golden_resnet18()
    golden_other_nets()
    golden_resnext()
    golden_augment()
    golden_sgd()
    golden_sampler()
