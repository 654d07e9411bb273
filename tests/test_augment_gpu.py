"""GPU parity of the input pipeline (main.py:30-35): the CUDA augmentation kernels against the numpy oracle and the
golden batch produced by torchvision itself (tests/golden/augment_seed123.npz). Bit-exact in fp32; the fused
im2col variant must equal the conv-stem operand built from those exact pixels."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_augment_matches_torchvision_golden_bit_exact():
    import b200cifar
    from b200cifar import data, ops
    g = np.load(os.path.join(G, "augment_seed123.npz"))
    n = g["images"].shape[0]
    ijf = data.draw_crop_flip(data.TorchCpuMT19937(int(g["seed"])), n)      # the product's own host draws
    imgs = torch.from_numpy(g["images"]).cuda()
    idx = torch.arange(n, device="cuda")
    out = torch.empty(n, 3, 32, 32, device="cuda")
    ops.augment_nchw(imgs, idx, torch.from_numpy(ijf).cuda(), out)
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), g["out"])                       # bit-exact with torchvision's batch


def test_augment_random_gather_and_im2col():
    import b200cifar
    from b200cifar import data, ops
    from oracle import augment as oaug
    images, _ = data.synthetic_cifar(512, seed=4)
    rng = data.TorchCpuMT19937(77)
    n = 200
    sel = data.randperm(rng, 512)[:n]
    ijf = data.draw_crop_flip(rng, n)
    ijf[0] = (0, 0, 0); ijf[1] = (8, 8, 1); ijf[2] = (4, 4, 0)               # corners and the identity crop
    dev = torch.from_numpy(images).cuda()
    out = torch.empty(n, 3, 32, 32, device="cuda")
    ops.augment_nchw(dev, torch.from_numpy(sel).cuda(), torch.from_numpy(ijf).cuda(), out)
    ref = oaug.augment(images[sel], ijf)
    assert np.array_equal(out.cpu().numpy(), ref)
    assert np.array_equal(out[2].cpu().numpy(), oaug.normalize_only(images[sel[2:3]])[0])    # transform_test
    col = torch.empty(n, 32, 32, 64, device="cuda", dtype=torch.bfloat16)
    ops.augment_im2col(dev, torch.from_numpy(sel).cuda(), torch.from_numpy(ijf).cuda(), col)
    col2 = torch.empty_like(col)
    ops.stem_im2col(out, col2)
    torch.cuda.synchronize()
    assert torch.equal(col, col2)
    # empty batch is a no-op, not an error
    ops.augment_nchw(dev, torch.zeros(0, dtype=torch.int64, device="cuda"),
                     torch.zeros(0, 3, dtype=torch.int8, device="cuda"), torch.empty(0, 3, 32, 32, device="cuda"))
