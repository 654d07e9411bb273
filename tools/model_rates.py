"""Training throughput of every supported network (BASELINE configs 2-5 + ResNet-50) on one GPU: whole step in a
CUDA graph, dataset resident in HBM, CUDA-event time over `steps` steps. Writes gpurun_out/model_rates.json."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import b200cifar
from b200cifar import data, engine, trainer

MODELS = [("resnet18", "resnet_spec", 1024), ("vgg16", "vgg_spec", 1024), ("preact_resnet18", "preact_resnet_spec", 1024),
          ("mobilenetv2", "mobilenetv2_spec", 512), ("resnet50", "resnet50_spec", 256), ("simpledla", "simple_dla_spec", 1024),
          ("mobilenet", "mobilenet_spec", 512), ("resnext29_2x64d", "resnext_spec", 512), ("resnext29_4x64d", "resnext29_4x64d_spec", 256),
          ("regnetx_200mf", "regnetx_spec", 512), ("regnetx_400mf", "regnetx_400mf_spec", 512),
          ("dla", "dla_spec", 1024), ("senet18", "senet_spec", 1024),
          ("regnety_400mf", "regnety_400mf_spec", 512), ("googlenet", "googlenet_spec", 256),
          ("preact_resnet50", "preact_resnet50_spec", 256)]
steps, warm = 30, 5
images, labels = data.synthetic_cifar(8192, seed=0)
out = []
for name, spec_fn, N in MODELS:
    if len(sys.argv) > 1 and name not in sys.argv[1:]:
        continue
    spec = getattr(engine, spec_fn)()
    params, buffers = engine.init_params(spec, 0)
    tr = trainer.Trainer(spec, N, lr=0.01)
    tr.load_params(params, buffers)
    tr.set_dataset(images, labels)
    rng = data.TorchCpuMT19937(3)
    idx = [data.randperm(rng, 8192)[:N] for _ in range(4)]
    ijf = [data.draw_crop_flip(rng, N).reshape(N, 3) for _ in range(4)]
    for k in range(warm):
        tr.step(idx[k % 4], ijf[k % 4])
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for k in range(steps):
        tr.step(idx[k % 4], ijf[k % 4])
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / steps
    rec = dict(model=name, batch=N, ms_per_step=round(ms, 3), images_per_s=round(N / ms * 1e3), loss=round(tr.read_loss(), 4))
    print(json.dumps(rec), flush=True)
    out.append(rec)
    del tr
    torch.cuda.empty_cache()
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/model_rates.json", "w"), indent=1)
