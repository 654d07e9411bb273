"""Debug: training-loss trajectory with the [tap][cin][cout] arena layout vs OIHW (same inputs)."""
import os, sys, json
sys.path.insert(0, ".")
import numpy as np, torch
mode = sys.argv[1]
os.environ["B2C_TCO_LAYOUT"] = mode
import b200cifar
from b200cifar import data, engine, trainer
from oracle import nets
N = 128
images, labels = data.learnable_cifar(4096 + 500, seed=1, noise=56.0)
params, buffers = nets.build_resnet_params(0)
tr = trainer.Trainer(engine.resnet_spec(), N, lr=0.05)
tr.load_params(params, {k: v for k, v in buffers.items() if not k.endswith("num_batches_tracked")})
tr.set_dataset(images[:4096], labels[:4096])
rng = data.TorchCpuMT19937(11)
losses = []
for ep in range(3):
    perm = data.randperm(data.TorchCpuMT19937(100 + ep), 4096)
    for b in data.batches(perm, N):
        tr.step(b, data.draw_crop_flip(rng, len(b)))
        losses.append(round(tr.read_loss(), 4))
    if len(sys.argv) > 2:
        print("eval", tr.evaluate(images[4096:], labels[4096:]))
print(mode, "mean(last 16) =", round(float(np.mean(losses[-16:])), 4), "mean(ep1) =", round(float(np.mean(losses[32:64])), 4), losses[::12])
g = tr.eng.named_grads()
print({k: float(v.norm()) for k, v in list(g.items())[:4]})
