"""images/s through the drop-in surface — accelerate(net) under the reference's loop body (main.py:101-107:
zero_grad / net(inputs) / criterion / backward / optimizer.step / loss.item) — next to the plain module on the same
GPU (torch / cuDNN, fp32 and bf16 autocast) and to trainer.Trainer (whole step in one CUDA graph).
Inputs are synthetic fp32 NCHW tensors already on the device, as a DataLoader + .to(device) would deliver them.
Writes gpurun_out/dropin_rates.json."""
import copy, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.nn as nn
import b200cifar
from b200cifar.accelerate import accelerate, FusedSGD
from oracle.stage_reference import reference_dir

ref = reference_dir()
if ref is not None:
    sys.path.insert(0, ref)
    import models.resnet as R
    make, src = R.ResNet18, "reference models/resnet.py"
else:
    from oracle import ref_modules
    make, src = ref_modules.ResNet18, "oracle/ref_modules.py"


def rate(net, opt, N, steps=30, warm=5, autocast=False, item=True):
    x = torch.randn(N, 3, 32, 32, device="cuda")
    t = torch.randint(0, 10, (N,), device="cuda")
    crit = nn.CrossEntropyLoss()
    def step():
        opt.zero_grad()
        if autocast:
            with torch.autocast("cuda", dtype=torch.bfloat16):
                loss = crit(net(x), t)
        else:
            loss = crit(net(x), t)
        loss.backward()
        opt.step()
        return loss.item() if item else None
    for _ in range(warm):
        step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    torch.cuda.synchronize()
    return N * steps / (time.perf_counter() - t0)


out = dict(module_source=src, rows=[])
for N in (128, 1024):
    torch.manual_seed(0)
    base = make().cuda()
    row = dict(batch=N)
    n1 = copy.deepcopy(base)
    row["plain_fp32_cudnn"] = rate(n1, torch.optim.SGD(n1.parameters(), lr=0.01, momentum=0.9, weight_decay=5e-4), N)
    n2 = copy.deepcopy(base)
    row["plain_bf16_autocast_cudnn"] = rate(n2, torch.optim.SGD(n2.parameters(), lr=0.01, momentum=0.9, weight_decay=5e-4), N, autocast=True)
    n3 = accelerate(copy.deepcopy(base))
    row["accelerate_torch_sgd"] = rate(n3, torch.optim.SGD(n3.parameters(), lr=0.01, momentum=0.9, weight_decay=5e-4), N)
    n4 = accelerate(copy.deepcopy(base))
    row["accelerate_fused_sgd"] = rate(n4, FusedSGD(n4, lr=0.01, momentum=0.9, weight_decay=5e-4), N)
    row["accelerate_fused_sgd_no_item"] = rate(n4, FusedSGD(n4, lr=0.01, momentum=0.9, weight_decay=5e-4), N, item=False)
    n5 = accelerate(copy.deepcopy(base), graphs=True)
    row["accelerate_graphs_torch_sgd"] = rate(n5, torch.optim.SGD(n5.parameters(), lr=0.01, momentum=0.9, weight_decay=5e-4), N)
    n6 = accelerate(copy.deepcopy(base), graphs=True)
    row["accelerate_graphs_fused_sgd"] = rate(n6, FusedSGD(n6, lr=0.01, momentum=0.9, weight_decay=5e-4), N)
    del n5, n6
    print(json.dumps(row), flush=True)
    out["rows"].append(row)
    del n1, n2, n3, n4
    torch.cuda.empty_cache()
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/dropin_rates.json", "w"), indent=1)
