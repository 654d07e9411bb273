"""GPU parity of the whole ResNet-18 training step (forward, loss, backward, SGD, BN buffers) against the
oracle on the same seeded inputs, through the C-ABI.

Two oracles, both pinned to the reference by tests/test_oracle.py:
  * fp32        oracle/nets.py (autograd). BASELINE.json north_star tolerance: per-layer activations and
                gradients within 1e-2 relative error (bf16 vs fp32). "Per layer" = each layer given the same
                inputs: that is tests/test_kernels_gpu.py (every kernel <= 1e-2 vs fp32 on identical inputs)
                and test_per_layer_teacher_forced below. END-TO-END through 20 layers the bf16 storage error
                compounds (~0.2-0.4 %/layer) and every ReLU mask that flips moves a gradient by O(1), so the
                end-to-end numbers vs fp32 are REPORTED (gpurun_out/parity_*.json) and bounded
                loosely here, next to what stock torch bf16 autocast shows on the same inputs.
  * bf16-store  oracle/nets_manual.py with round_bf16=True: rounds exactly where the CUDA path stores bf16.
                Fed the same tensors block by block (teacher forcing) the CUDA path must match it tightly,
                forward and backward; this is the test that catches kernel bugs. End to end two bf16
                pipelines decorrelate after a few layers, so that comparison is only loosely bounded.
Relative error = ||a-b||_F / ||b||_F per tensor.
"""
import json
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

TOL = 1e-2


def rel(a, b):
    a, b = a.double(), b.double()
    return ((a - b).norm() / b.norm().clamp_min(1e-30)).item()


def dump(name, obj):
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(obj, open(os.path.join("gpurun_out", name), "w"), indent=1)


@pytest.fixture(scope="module")
def setup():
    import b200cifar
    from b200cifar import engine, ops
    from oracle import nets, nets_manual
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    N = 128                                              # BASELINE config 1 / per-GPU batch of config 2 at 8 GPUs
    params, buffers = nets.build_resnet_params(0)
    rs = np.random.RandomState(1)
    x = torch.from_numpy(rs.standard_normal((N, 3, 32, 32)).astype(np.float32)).cuda()
    target = torch.from_numpy(rs.randint(0, 10, size=(N,)).astype(np.int64)).cuda()
    cp = {k: v.cuda() for k, v in params.items()}
    fb = {k: v.cuda() for k, v in buffers.items()}
    taps32 = {}
    loss32, logits32, grads32 = nets.train_step_grads(x, target, cp, fb, taps=taps32)
    eb = {k: v.cuda() for k, v in buffers.items()}
    tapsE = {}
    with torch.no_grad():
        lossE, logitsE, gradsE = nets_manual.resnet_train_step(x, target, cp, eb, True, taps=tapsE)
    eng = engine.Engine(engine.resnet_spec(), N)
    eng.load_params(params, {k: v for k, v in buffers.items() if not k.endswith("num_batches_tracked")})
    ops.stem_im2col(x, eng.x0)
    eng.forward(True)
    loss = torch.zeros(1, device="cuda")
    ops.ce_loss(eng.logits, target, N, 10, eng.dlogits, loss)
    eng.zero_grad()
    eng.backward()
    torch.cuda.synchronize()
    return dict(eng=eng, ops=ops, taps32=taps32, tapsE=tapsE, loss32=loss32, logits32=logits32, grads32=grads32,
                lossE=lossE, logitsE=logitsE, gradsE=gradsE, loss=loss, fb=fb, eb=eb, cp=cp, x=x, target=target,
                params=params, buffers=buffers, N=N)


def test_forward_vs_bf16_storage_oracle(setup):
    """End to end the two bf16 pipelines decorrelate (a 1e-5 accumulation-order difference flips bf16
    roundings, and flipped roundings feed the next layer), so only the first layers match tightly; deeper
    layers are bounded by the storage-noise level and covered tightly by the teacher-forced tests below."""
    eng, taps = setup["eng"], setup["tapsE"]
    errs = {u.name: rel(u.y.float().permute(0, 3, 1, 2), taps[u.name]) for u in eng.units if u.name in taps}
    errs["logits"] = rel(eng.logits, setup["logitsE"])
    errs["loss_abs"] = abs(setup["loss"].item() - setup["lossE"].item())
    dump("parity_forward_vs_bf16store.json", errs)
    assert errs["conv1"] < 1e-4 and errs["layer1.0.conv1"] < 1e-3 and errs["layer1.0.conv2"] < 2e-3
    assert max(v for k, v in errs.items() if k != "loss_abs") < 4e-2, errs
    assert errs["loss_abs"] < 5e-3


def _nhwc(t):
    return t.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16)


@pytest.mark.parametrize("N", [128, 1024])
def test_backward_teacher_forced_per_block(setup, N):
    """Every backward op of the plan (head, 8 BasicBlocks, stem) run on the bf16-storage oracle's own tensors
    for that block: parameter gradients and the input gradient must match the oracle tightly.
    N = 128 is BASELINE config 1 (and config 2 per GPU at 8 GPUs); N = 1024 is the batch bench.py is quoted on
    (config 2 on one GPU): other tile counts, split-K depths (wgrad K = 1 048 576 on layer 1), scratch use and
    persistent-loop wave counts than the small case."""
    from b200cifar import engine, ops
    from oracle import nets_manual
    cp = setup["cp"]
    if N == setup["N"]:
        x, target = setup["x"], setup["target"]
    else:
        rs = np.random.RandomState(1000 + N)
        x = torch.from_numpy(rs.standard_normal((N, 3, 32, 32)).astype(np.float32)).cuda()
        target = torch.from_numpy(rs.randint(0, 10, size=(N,)).astype(np.int64)).cuda()
    eb = {k: v.clone().cuda() for k, v in setup["buffers"].items()}
    trace = {}
    with torch.no_grad():
        _, _, gradsE = nets_manual.resnet_train_step(x, target, cp, eb, True, trace=trace)
    eng = engine.Engine(engine.resnet_spec(), N)
    eng.load_params(setup["params"], {k: v for k, v in setup["buffers"].items()
                                      if not k.endswith("num_batches_tracked")})

    def force_unit(u, ref):
        u.y.copy_(_nhwc(ref.y)); u.mean.copy_(ref.mean); u.rstd.copy_(ref.rstd)
        # the forward's scale/shift (what bn_finalize would have left): plain BN+ReLU units recompute their mask
        # from y with them
        u.scale[:u.gamma.numel()].copy_(u.gamma * u.rstd[:u.gamma.numel()])
        u.shift[:u.gamma.numel()].copy_(u.beta - u.mean[:u.gamma.numel()] * u.scale[:u.gamma.numel()])

    # forward state of the whole net from the oracle
    ops.stem_im2col(x, eng.x0)
    force_unit(eng.stem_unit, trace["stem"]); eng.a0.copy_(_nhwc(trace["a0"]))
    for bb, (c1, a1, c2, cs, out) in zip(reversed(eng.block_bufs), trace["blocks"]):
        force_unit(bb["c1"], c1); bb["a1"].copy_(_nhwc(a1)); force_unit(bb["c2"], c2); bb["out"].copy_(_nhwc(out))
        # the ReLU bit mask the forward apply pass would have written next to `out` (bit e of byte i = out[8i+e] > 0)
        w8 = (1 << torch.arange(8, device="cuda")).to(torch.int32)
        bb["bits"].copy_(((bb["out"].view(-1, 8) > 0).to(torch.int32) * w8).sum(1).to(torch.uint8))
        if cs is not None:
            force_unit(bb["cs"], cs)
    eng.pooled.copy_(trace["pooled"]); eng.dlogits.copy_(trace["dlogits"])
    eng.zero_grad()
    errs = {}
    eng.bwd_ops[0]()                                                     # head
    torch.cuda.synchronize()
    errs["head.dout"] = rel(eng.block_bufs[0]["dout"].float().permute(0, 3, 1, 2), trace["dout_in"][0])
    for i, bb in enumerate(eng.block_bufs):                              # blocks, last first
        bb["dout"].copy_(_nhwc(trace["dout_in"][i]))                     # teacher forcing
        eng.bwd_ops[1 + i]()
        torch.cuda.synchronize()
        errs[bb["c1"].name.rsplit(".", 1)[0] + ".dx"] = rel(bb["dx"].float().permute(0, 3, 1, 2), trace["dx_out"][i])
    eng.block_bufs[-1]["dx"].copy_(_nhwc(trace["dx_out"][-1]))
    eng.bwd_ops[-1]()                                                    # stem
    torch.cuda.synchronize()
    for name, g in eng.named_grads().items():
        if gradsE[name].norm().item() > 1e-7:
            errs[name] = rel(g, gradsE[name])
    dump(f"parity_backward_teacher_forced_n{N}.json", errs)
    assert max(errs.values()) < 5e-3, sorted(errs.items(), key=lambda kv: -kv[1])[:8]
    del trace, eng
    torch.cuda.empty_cache()


def test_bn_running_stats(setup):
    eng, eb, fb = setup["eng"], setup["eb"], setup["fb"]
    for name, b in eng.named_buffers().items():
        assert rel(b, eb[name]) < 2e-2, name
        if name.endswith("running_mean"):       # vs fp32: error of the mean measured in units of the std
            var = fb[name.replace("running_mean", "running_var")]
            assert ((b - fb[name]).abs() / (10 * var - 9).clamp_min(1e-6).sqrt() / 0.1).max().item() < 1e-2, name
        else:
            assert rel(b, fb[name]) < 1e-2, name


def test_end_to_end_vs_fp32_reported(setup):
    """End-to-end bf16-vs-fp32 differences: reported, loosely bounded, and compared with stock torch bf16."""
    from oracle import nets
    eng, taps32 = setup["eng"], setup["taps32"]
    fwd = {u.name: rel(u.y.float().permute(0, 3, 1, 2), taps32[u.name].detach()) for u in eng.units
           if u.name in taps32}
    fwd["logits"] = rel(eng.logits, setup["logits32"])
    fwd["loss_abs"] = abs(setup["loss"].item() - setup["loss32"].item())
    grads = {n: rel(g, setup["grads32"][n]) for n, g in eng.named_grads().items()
             if setup["grads32"][n].norm().item() > 1e-7}
    # stock torch bf16 autocast on the same inputs, for context
    leaves = {k: v.detach().clone().requires_grad_(True) for k, v in setup["cp"].items()}
    ab = {k: v.clone().cuda() for k, v in setup["buffers"].items()}
    with torch.autocast("cuda", dtype=torch.bfloat16):
        lg = nets.resnet_forward(setup["x"], leaves, ab, training=True)
        ls = F.cross_entropy(lg.float(), setup["target"])
    ag = torch.autograd.grad(ls, list(leaves.values()))
    auto = {n: rel(g.float(), setup["grads32"][n]) for n, g in zip(leaves.keys(), ag)
            if setup["grads32"][n].norm().item() > 1e-7}
    dump("parity_vs_fp32.json", dict(forward=fwd, grads=grads, torch_autocast_grads=auto,
                                     torch_autocast_logits=rel(lg.float(), setup["logits32"])))
    assert fwd["conv1"] < TOL and fwd["layer1.0.conv1"] < TOL          # first layers: no compounding yet
    assert fwd["logits"] < 5e-2 and fwd["loss_abs"] < 1e-2
    med_ours = float(np.median(list(grads.values())))
    med_auto = float(np.median(list(auto.values())))
    assert med_ours < 1.5 * med_auto + 1e-2, (med_ours, med_auto)      # no worse than stock bf16 training


def test_per_layer_teacher_forced(setup):
    """Each of the 20 convs of ResNet-18 (stem, conv1 / conv2 of the 8 blocks, the 3 shortcut convs) fed the fp32
    oracle's own input for that layer: output within 1e-2 of the fp32 oracle (north_star's per-layer bound)."""
    from b200cifar import ops
    eng, taps32, N = setup["eng"], setup["taps32"], setup["N"]
    errs = {}
    for u in eng.units:
        y = torch.empty_like(u.y)
        if u.stem:                                         # models/resnet.py:132: the network input itself
            col = torch.empty_like(eng.x0)
            ops.stem_im2col(setup["x"], col)
            ops.conv_fprop(u.desc, col, u.wf, y)
        else:
            blk = ".".join(u.name.split(".")[:2])          # "layerL.B"
            if u.name.endswith("conv2"):
                src = blk + ".a1"                          # relu(bn1(conv1(x))), resnet.py:47
            else:                                          # conv1 and shortcut.0 read the block input, resnet.py:47,49
                li, bi = int(blk[5]), int(blk[7])
                src = "act0" if blk == "layer1.0" else (f"layer{li}.{bi - 1}.out" if bi > 0 else f"layer{li - 1}.1.out")
            xb = taps32[src].detach().permute(0, 2, 3, 1).contiguous().to(torch.bfloat16)
            ops.conv_fprop(ops.conv_desc(N, u.H, u.W, u.cin, u.cout, u.k, u.stride), xb, u.wf, y)
        torch.cuda.synchronize()
        errs[u.name] = rel(y.float().permute(0, 3, 1, 2), taps32[u.name].detach())
    dump("parity_teacher_forced.json", errs)
    assert len(errs) == 20 and max(errs.values()) < TOL, errs


def test_per_layer_gradients_teacher_forced_vs_fp32(setup):
    """The backward counterpart against fp32: every conv's data gradient and weight gradient computed from the
    fp32 oracle's own dy / x for that layer (autograd, nets.py) within 1e-2 of the fp32 result."""
    from b200cifar import ops
    eng, N, cp = setup["eng"], setup["N"], setup["cp"]
    errs = {}
    g = torch.Generator(device="cuda").manual_seed(77)
    for u in eng.units:
        if u.stem:
            continue
        w = cp[u.name + ".weight"]
        x32 = torch.randn(N, u.cin_true, u.H, u.W, device="cuda", generator=g)
        dy32 = torch.randn(N, u.cout_true, u.Ho, u.Wo, device="cuda", generator=g)
        xb = x32.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16)
        dyb = dy32.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16)
        dx = torch.zeros(N, u.H, u.W, u.cin, device="cuda", dtype=torch.bfloat16)
        ops.conv_dgrad(ops.conv_desc(N, u.H, u.W, u.cin, u.cout, u.k, u.stride), dyb, u.wd, dx)
        dw = torch.zeros(u.cout_true, u.cin_true, u.k, u.k, device="cuda")
        ops.conv_wgrad(ops.conv_desc(N, u.H, u.W, u.cin, u.cout, u.k, u.stride), xb, dyb, dw)
        torch.cuda.synchronize()
        pad = u.k // 2
        ref_dx = torch.nn.grad.conv2d_input(x32.shape, w, dy32, stride=u.stride, padding=pad)
        ref_dw = torch.nn.grad.conv2d_weight(x32, w.shape, dy32, stride=u.stride, padding=pad)
        errs[u.name + ".dx"] = rel(dx.float().permute(0, 3, 1, 2), ref_dx)
        errs[u.name + ".dw"] = rel(dw, ref_dw)
    dump("parity_gradients_teacher_forced_fp32.json", errs)
    assert len(errs) == 38 and max(errs.values()) < TOL, sorted(errs.items(), key=lambda kv: -kv[1])[:6]


def test_sgd_trajectory(setup):
    """4 steps of forward/backward/SGD vs the bf16-storage oracle driven by the same update rule."""
    from b200cifar import engine, ops
    from oracle import nets_manual
    N, x, target = setup["N"], setup["x"], setup["target"]
    params, buffers = setup["params"], setup["buffers"]
    eng = engine.Engine(engine.resnet_spec(), N)
    eng.load_params(params, {k: v for k, v in buffers.items() if not k.endswith("num_batches_tracked")})
    lr, mom, wd = 0.02, 0.9, 5e-4
    hyper = torch.tensor([lr, mom, wd, 1.0, 1.0], device="cuda")
    P = {k: v.clone().cuda() for k, v in params.items()}
    M = {k: torch.zeros_like(v) for k, v in P.items()}
    cb = {k: v.clone().cuda() for k, v in buffers.items()}
    loss = torch.zeros(1, device="cuda")
    for step in range(4):
        with torch.no_grad():
            l_ref, _, g_ref = nets_manual.resnet_train_step(x, target, P, cb, True)
            for k in P:                                   # optim.SGD, main.py:87-88
                d = g_ref[k] + wd * P[k]
                M[k] = d.clone() if step == 0 else mom * M[k] + d
                P[k] = P[k] - lr * M[k]
        ops.stem_im2col(x, eng.x0)
        eng.zero_grad()
        eng.forward(True)
        ops.ce_loss(eng.logits, target, N, 10, eng.dlogits, loss)
        eng.backward()
        hyper[4] = 1.0 if step == 0 else 0.0
        ops.sgd_step(eng.P, eng.G, eng.M, hyper)
        eng.repack()
        torch.cuda.synchronize()
        assert abs(loss.item() - l_ref.item()) < 2e-2, (step, loss.item(), l_ref.item())
    errs = {name: rel(p, P[name]) for name, p in eng.named_params().items() if P[name].norm() > 1e-2}
    dump("parity_sgd_traj.json", errs)
    assert max(errs.values()) < 3e-2, sorted(errs.items(), key=lambda kv: -kv[1])[:5]   # chaotic divergence of two bf16 runs
