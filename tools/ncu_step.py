"""One training step for Nsight Compute: builds the batch-1024 ResNet-18 trainer (or --model), runs warm-up steps with
profiling off, then exactly ONE eager step between cudaProfilerStart/Stop. Use with

    ncu --profile-from-start off --set full --clock-control none --import-source on \
        -k regex:'conv_halo2|tap_gemm|wgrad' -o gpurun_out/r2_conv_step python tools/ncu_step.py

(never a bench: numbers printed under a profiler are not measurements)."""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200cifar
from b200cifar import data, engine, trainer

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="resnet_spec")
ap.add_argument("--batch", type=int, default=1024)
a = ap.parse_args()
spec = getattr(engine, a.model)()
params, bufs = engine.init_params(spec, 0)
tr = trainer.Trainer(spec, a.batch, use_graph=False)
tr.load_params(params, bufs)
images, labels = data.synthetic_cifar(4096, seed=0)
tr.set_dataset(images, labels)
rng = data.TorchCpuMT19937(1)
idx = data.randperm(rng, 4096)[:a.batch] if a.batch <= 4096 else None
ijf = data.draw_crop_flip(rng, a.batch)
for _ in range(3):
    tr.step(idx, ijf)
torch.cuda.synchronize()
torch.cuda.profiler.start()
tr.step(idx, ijf)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("loss", tr.read_loss())
