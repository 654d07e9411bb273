"""GPU parity of the whole ResNet-18 training step (forward, loss, backward, SGD, BN buffers) against the
fp32 oracle (oracle/nets.py, pinned to the reference by tests/test_oracle.py) on the same seeded inputs.

Tolerance (BASELINE.json north_star): per-layer forward activations and gradients within 1e-2 relative
error (bf16 storage, fp32 accumulation) versus fp32. Relative error = ||a-b||_F / ||b||_F per tensor.
"""
import json
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 1e-2


def rel(a, b):
    a, b = a.double(), b.double()
    return ((a - b).norm() / b.norm().clamp_min(1e-30)).item()


@pytest.fixture(scope="module")
def setup():
    import b200cifar
    from b200cifar import engine, ops
    from oracle import nets
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    N = 128                                              # BASELINE config 1 / per-GPU batch of config 2 at 8 GPUs
    params, buffers = nets.build_resnet_params(0)
    rs = np.random.RandomState(1)
    x = torch.from_numpy(rs.standard_normal((N, 3, 32, 32)).astype(np.float32)).cuda()
    target = torch.from_numpy(rs.randint(0, 10, size=(N,)).astype(np.int64)).cuda()
    # oracle on the GPU in fp32 (TF32 off) — same arithmetic as the CPU oracle, just faster
    cp = {k: v.cuda() for k, v in params.items()}
    cb = {k: v.cuda() for k, v in buffers.items()}
    taps = {}
    loss_ref, logits_ref, grads_ref = nets.train_step_grads(x, target, cp, cb, taps=taps)
    eng = engine.Engine(engine.resnet_spec(), N)
    eng.load_params(params, {k: v for k, v in buffers.items() if not k.endswith("num_batches_tracked")})
    ops.stem_im2col(x, eng.x0)
    eng.forward(True)
    loss = torch.zeros(1, device="cuda")
    ops.ce_loss(eng.logits, target, N, 10, eng.dlogits, loss)
    eng.zero_grad()
    eng.backward()
    torch.cuda.synchronize()
    return dict(eng=eng, ops=ops, taps=taps, loss_ref=loss_ref, logits_ref=logits_ref, grads_ref=grads_ref,
                loss=loss, cb=cb, cp=cp, x=x, target=target, params=params, buffers=buffers, N=N)


def test_forward_activations(setup):
    eng, taps = setup["eng"], setup["taps"]
    errs = {}
    for u in eng.units:
        ref = taps.get(u.name)
        if ref is None:
            continue
        errs[u.name] = rel(u.y.float().permute(0, 3, 1, 2), ref.detach())
    errs["logits"] = rel(eng.logits, setup["logits_ref"])
    errs["loss_abs"] = abs(setup["loss"].item() - setup["loss_ref"].item())
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(errs, open("gpurun_out/parity_forward.json", "w"), indent=1)
    worst = max(v for k, v in errs.items() if k != "loss_abs")
    assert worst < TOL, errs
    assert errs["loss_abs"] < 1e-2


def test_gradients(setup):
    eng, grads_ref = setup["eng"], setup["grads_ref"]
    errs = {}
    for name, g in eng.named_grads().items():
        ref = grads_ref[name]
        if name.endswith("bias") and ref.norm().item() < 1e-6:
            continue
        errs[name] = rel(g, ref)
    json.dump(errs, open("gpurun_out/parity_grads.json", "w"), indent=1)
    worst = max(errs.values())
    assert worst < TOL, sorted(errs.items(), key=lambda kv: -kv[1])[:8]


def test_bn_running_stats(setup):
    eng, cb = setup["eng"], setup["cb"]
    for name, b in eng.named_buffers().items():
        assert rel(b, cb[name]) < 2e-3, name


def test_sgd_trajectory(setup):
    """3 steps of forward/backward/SGD vs the oracle driven by torch.optim.SGD on the fp32 restatement."""
    from b200cifar import engine, ops
    from oracle import nets
    N, x, target = setup["N"], setup["x"], setup["target"]
    params, buffers = setup["params"], setup["buffers"]
    eng = engine.Engine(engine.resnet_spec(), N)
    eng.load_params(params, {k: v for k, v in buffers.items() if not k.endswith("num_batches_tracked")})
    hyper = torch.tensor([0.1, 0.9, 5e-4, 1.0, 1.0], device="cuda")
    leaves = {k: v.detach().clone().cuda().requires_grad_(True) for k, v in params.items()}
    cb = {k: v.clone().cuda() for k, v in buffers.items()}
    opt = torch.optim.SGD(list(leaves.values()), lr=0.1, momentum=0.9, weight_decay=5e-4)
    loss = torch.zeros(1, device="cuda")
    for step in range(3):
        opt.zero_grad()
        l_ref = torch.nn.functional.cross_entropy(nets.resnet_forward(x, leaves, cb, training=True), target)
        l_ref.backward()
        opt.step()
        ops.stem_im2col(x, eng.x0)
        eng.zero_grad()
        eng.forward(True)
        ops.ce_loss(eng.logits, target, N, 10, eng.dlogits, loss)
        eng.backward()
        hyper[4] = 1.0 if step == 0 else 0.0
        ops.sgd_step(eng.P, eng.G, eng.M, hyper)
        eng.repack()
        torch.cuda.synchronize()
        assert abs(loss.item() - l_ref.item()) < 3e-2, (step, loss.item(), l_ref.item())
    worst = max(rel(p, leaves[name].detach()) for name, p in eng.named_params().items()
                if leaves[name].detach().norm() > 1e-3)
    assert worst < TOL
