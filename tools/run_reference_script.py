#!/usr/bin/env python
"""Run the reference's OWN main.py / main_dist.py — the unmodified files — for a few steps on synthetic data,
with or without the B200 path under their modules, and record what they computed.

    python tools/run_reference_script.py --script main.py --model resnet.ResNet18 --out a.json
    python tools/run_reference_script.py --script main.py --model resnet.ResNet18 --plain --out b.json
    python tools/run_reference_script.py --script main_dist.py --model resnet.ResNet18 --out c.json -- \
        --dist --world_size 1 --rank 0 --dist_url tcp://127.0.0.1:29517 --batch_size 128 --epochs 1 \
        --output_dir OUT --workers 2

The script file is executed as it is (runpy, run_name "__main__") from the reference tree (/root/reference, or the
copy staged by oracle/stage_reference.py on machines without it). The harness only prepares the process around it;
each item names the line of the reference that makes it necessary:
  * a `stty` stub on PATH — utils.py:46-47 reads the terminal size at import and crashes without a tty;
  * `torchvision.datasets.CIFAR10` (main.py:42-47, main_dist.py:105-119) replaced by a small synthetic learnable
    set with the same item protocol (PIL image through `transform`) — the real files are md5-checked downloads;
  * the model: main.py:71 hard-codes `SimpleDLA()`, main_dist.py:136 `models.ResNet152(args.amp)`; the harness
    re-binds that NAME in the `models` package to a constructor of --model (the scripts select models by editing
    that line);
  * `net = accelerate(net)`: applied the moment the script moves the network to the GPU (`net.to(device)`,
    main.py:72; `net.cuda()`, main_dist.py:137), unless --plain;
  * `attach_ddp()` on the DistributedDataParallel the script builds (main_dist.py:141-144), unless --no-attach;
  * one visible GPU: CUDA_VISIBLE_DEVICES=0, so main.py:73-75's nn.DataParallel has a single device and
    main_dist.py:58's mp.spawn has one worker, which is run in-process (the spawn start method would re-import
    the script in a child without any of this set-up);
  * an end: main.py loops for 200 epochs; the harness stops it after --stop-epochs `scheduler.step()` calls.
Recorded: every criterion() value (train and test), wall time of the training loop, the checkpoint file the script
wrote ({'net','acc','epoch'} with `module.`-prefixed keys, main.py:138-147) re-loaded into a plain fp32 copy of the
model and evaluated on the same test set.
"""
import argparse
import json
import os
import runpy
import stat
import sys
import tempfile
import time

os.environ.setdefault("CUDA_VISIBLE_DEVICES", "0")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


class _Stop(Exception):
    pass


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--script", default="main.py", choices=["main.py", "main_dist.py"])
    ap.add_argument("--model", default="resnet.ResNet18", help="<file in models/>.<constructor>, e.g. vgg.VGG:VGG16")
    ap.add_argument("--plain", action="store_true", help="no accelerate(): the reference's own torch/cuDNN path, fp32")
    ap.add_argument("--no-attach", action="store_true", help="leave DDP's Reducer to do the all-reduce itself")
    ap.add_argument("--train", type=int, default=428, help="synthetic training samples (428 = 3 x 128 + 44)")
    ap.add_argument("--test", type=int, default=200)
    ap.add_argument("--stop-epochs", type=int, default=1)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--out", default=None)
    ap.add_argument("rest", nargs=argparse.REMAINDER, help="arguments of the script itself, after --")
    args = ap.parse_args()
    script_args = args.rest[1:] if args.rest[:1] == ["--"] else args.rest

    from oracle.stage_reference import reference_dir
    ref = reference_dir()
    if ref is None:
        raise SystemExit("reference files not found: run `python oracle/stage_reference.py` where /root/reference exists")

    # ---- stty stub (utils.py:46-47)
    shim = tempfile.mkdtemp(prefix="b2c_shim_")
    with open(os.path.join(shim, "stty"), "w") as f:
        f.write("#!/bin/sh\necho 24 80\n")
    os.chmod(os.path.join(shim, "stty"), stat.S_IRWXU)
    os.environ["PATH"] = shim + os.pathsep + os.environ.get("PATH", "")
    workdir = tempfile.mkdtemp(prefix="b2c_run_")
    os.chdir(workdir)                                    # ./data, ./checkpoint land here
    sys.path.insert(0, ref)

    import numpy as np
    import torch
    import torchvision
    from PIL import Image

    import b200cifar
    from b200cifar import data as hostdata
    from b200cifar.accelerate import accelerate, attach_ddp

    torch.backends.cudnn.allow_tf32 = False              # the plain arm is the fp32 oracle
    torch.backends.cuda.matmul.allow_tf32 = False

    # ---- dataset stand-in (same item protocol as torchvision/datasets/cifar.py:104-126)
    class SyntheticCIFAR10(torch.utils.data.Dataset):
        def __init__(self, root=None, train=True, transform=None, target_transform=None, download=False):
            n = args.train if train else args.test
            self.data, labels = hostdata.learnable_cifar(n, seed=11 if train else 12, noise=40.0)
            self.targets = labels.tolist()
            self.transform, self.target_transform = transform, target_transform

        def __getitem__(self, i):
            img, target = Image.fromarray(self.data[i]), self.targets[i]
            if self.transform is not None:
                img = self.transform(img)
            if self.target_transform is not None:
                target = self.target_transform(target)
            return img, target

        def __len__(self):
            return len(self.data)
    torchvision.datasets.CIFAR10 = SyntheticCIFAR10

    # ---- model selection + accelerate() at the script's own .to(device) / .cuda()
    import models
    mod_name, ctor_name = args.model.split(".", 1)
    ctor_name, _, ctor_arg = ctor_name.partition(":")
    ctor = getattr(getattr(models, mod_name), ctor_name)
    built = {}

    def make_plain():
        return ctor(ctor_arg) if ctor_arg else ctor()

    def factory(*_a, **_k):                              # ResNet152(args.amp): amp stays False (fp16 AMP is not built)
        net = make_plain()
        built["net"] = net
        if not args.plain:
            for name in ("to", "cuda"):
                orig = getattr(net, name)

                def moved(*a, _orig=orig, **k):
                    r = _orig(*a, **k)
                    if next(r.parameters()).is_cuda and not hasattr(r, "_b2c_state"):
                        accelerate(r)
                    return r
                setattr(net, name, moved)
        return net
    models.SimpleDLA = factory                           # main.py:71
    models.ResNet152 = factory                           # main_dist.py:136

    # ---- DDP: attach the arena-bucket all-reduce to the wrapper the script constructs (main_dist.py:141-144)
    DDP = torch.nn.parallel.DistributedDataParallel

    class AttachedDDP(DDP):
        def __init__(self, module, *a, **k):
            super().__init__(module, *a, **k)
            if hasattr(module, "_b2c_state") and not args.no_attach:
                attach_ddp(self)
                built["attached"] = True
    torch.nn.parallel.DistributedDataParallel = AttachedDDP

    # ---- mp.spawn with one visible GPU: run the single worker in this process (main_dist.py:58)
    import torch.multiprocessing as mp

    def spawn_inline(fn, args=(), nprocs=1, **_k):
        if nprocs != 1:
            raise RuntimeError("harness: one visible GPU expected (CUDA_VISIBLE_DEVICES=0)")
        fn(0, *args)
    mp.spawn = spawn_inline

    # ---- records: criterion values, optimizer steps, loop wall time
    rec = dict(train_losses=[], test_losses=[], sgd_steps=0, t_first_step=None, t_last_step=None)
    ce_forward = torch.nn.CrossEntropyLoss.forward

    def ce_recording(self, inp, target):
        out = ce_forward(self, inp, target)
        (rec["train_losses"] if torch.is_grad_enabled() else rec["test_losses"]).append(float(out.item()))
        return out
    torch.nn.CrossEntropyLoss.forward = ce_recording
    sgd_step = torch.optim.SGD.step

    def sgd_recording(self, *a, **k):
        r = sgd_step(self, *a, **k)
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        now = time.perf_counter()
        rec["sgd_steps"] += 1
        rec["t_first_step"] = rec["t_first_step"] or now
        rec["t_last_step"] = now
        return r
    torch.optim.SGD.step = sgd_recording

    # ---- an end for main.py's 200-epoch loop: stop after --stop-epochs scheduler.step() calls (main.py:154)
    sched_step = torch.optim.lr_scheduler.CosineAnnealingLR.step

    def sched_counting(self, *a, **k):
        r = sched_step(self, *a, **k)
        if self._step_count > args.stop_epochs:          # the constructor's own initial step() leaves _step_count at 1
            raise _Stop()
        return r
    torch.optim.lr_scheduler.CosineAnnealingLR.step = sched_counting

    torch.manual_seed(args.seed)
    np.random.seed(args.seed)
    sys.argv = [os.path.join(ref, args.script)] + script_args
    t0 = time.perf_counter()
    try:
        runpy.run_path(os.path.join(ref, args.script), run_name="__main__")
    except _Stop:
        pass
    wall = time.perf_counter() - t0
    torch.nn.CrossEntropyLoss.forward = ce_forward
    torch.optim.SGD.step = sgd_step

    net = built.get("net")
    out = dict(script=args.script, model=args.model, accelerated=not args.plain, reference_dir=ref,
               ddp_attached=bool(built.get("attached")), train_losses=rec["train_losses"],
               test_losses=rec["test_losses"], sgd_steps=rec["sgd_steps"], wall_s=wall,
               kernels_used=bool(net is not None and hasattr(net, "_b2c_state")))
    if rec["sgd_steps"] > 1:
        out["train_loop_s_after_first_step"] = rec["t_last_step"] - rec["t_first_step"]

    # ---- the checkpoint file the script wrote (main.py:138-147 / main_dist.py:243-250)
    ckpt_path = None
    for cand in ("./checkpoint/ckpt.pth",) + tuple(os.path.join(a, "ckpt.pth") for a in script_args if os.path.isdir(a)):
        if os.path.isfile(cand):
            ckpt_path = cand
    if ckpt_path is not None:
        dev = "cuda" if torch.cuda.is_available() else "cpu"
        ck = torch.load(ckpt_path, map_location=dev)
        keys = list(ck["net"].keys())
        plain = make_plain().to(dev)
        if keys[0].startswith("module."):                # saved through the DP / DDP wrapper, as on a GPU
            plain = torch.nn.DataParallel(plain, device_ids=[0])
        missing = plain.load_state_dict(ck["net"])
        plain.eval()
        ds = SyntheticCIFAR10(train=False)
        mean = torch.tensor([0.4914, 0.4822, 0.4465], device=dev).view(1, 3, 1, 1)
        std = torch.tensor([0.2023, 0.1994, 0.2010], device=dev).view(1, 3, 1, 1)
        x = (torch.from_numpy(ds.data).to(dev).permute(0, 3, 1, 2).float() / 255 - mean) / std
        t = torch.tensor(ds.targets, device=dev)
        with torch.no_grad():
            logits = torch.cat([plain(x[i:i + 100]) for i in range(0, len(x), 100)])
        out["ckpt"] = dict(path=ckpt_path, top_level_keys=sorted(ck.keys()), n_entries=len(keys),
                           module_prefix=all(k.startswith("module.") for k in keys), acc=float(ck["acc"]),
                           epoch=int(ck["epoch"]), load_result=str(missing),
                           plain_eval_loss=float(torch.nn.functional.cross_entropy(logits, t).item()),
                           plain_eval_acc=float((logits.argmax(1) == t).float().mean().item() * 100))
    if torch.distributed.is_initialized():
        torch.distributed.destroy_process_group()
    text = json.dumps(out)
    if args.out:
        with open(os.path.join(ROOT, args.out) if not os.path.isabs(args.out) else args.out, "w") as f:
            f.write(text + "\n")
    print("\n" + text)


if __name__ == "__main__":
    main()
