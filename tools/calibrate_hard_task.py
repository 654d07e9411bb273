"""Calibration aid for tests/test_accuracy_curve_gpu.py: final top-1 of the fp32 oracle (torch ops, TF32 off) on
data.hard_cifar for a few prototype amplitudes — pick the one that ends at 80-90 %. Writes
gpurun_out/hard_task_calibration.json."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.nn.functional as F
import b200cifar
from b200cifar import data
from oracle import augment as oaug
from oracle import nets

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
N, epochs, lr0 = 128, 6, 0.02
out = {}
for amp in [float(a) for a in (sys.argv[1:] or ["3", "4.5", "6", "9"])]:
    tr_x, tr_y = data.hard_cifar(8192, seed=1000, amplitude=amp)
    te_x, te_y = data.hard_cifar(4000, seed=1001, amplitude=amp)
    params, buffers = nets.build_resnet_params(0)
    leaves = {k: v.detach().clone().cuda().requires_grad_(True) for k, v in params.items()}
    cb = {k: v.clone().cuda() for k, v in buffers.items()}
    opt = torch.optim.SGD(list(leaves.values()), lr=lr0, momentum=0.9, weight_decay=5e-4)
    rng = data.TorchCpuMT19937(11)
    curve = []
    for ep in range(epochs):
        opt.param_groups[0]["lr"] = 0.5 * lr0 * (1 + np.cos(np.pi * ep / epochs))
        perm = data.randperm(data.TorchCpuMT19937(100 + ep), len(tr_y))
        for b in data.batches(perm, N):
            ijf = data.draw_crop_flip(rng, len(b))
            xb = torch.from_numpy(oaug.augment(tr_x[b], ijf)).cuda()
            tb = torch.from_numpy(tr_y[b]).cuda()
            opt.zero_grad()
            F.cross_entropy(nets.resnet_forward(xb, leaves, cb, training=True), tb).backward()
            opt.step()
        hits = 0
        with torch.no_grad():
            for s in range(0, len(te_y), 200):
                xe = torch.from_numpy(oaug.normalize_only(te_x[s:s + 200])).cuda()
                hits += (nets.resnet_forward(xe, leaves, cb, training=False).argmax(1).cpu().numpy() == te_y[s:s + 200]).sum()
        curve.append(hits / len(te_y))
    out[str(amp)] = curve
    print(amp, curve, flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/hard_task_calibration.json", "w"))
