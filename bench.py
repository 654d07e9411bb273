#!/usr/bin/env python
"""bench.py — CIFAR-10 training throughput of the B200-native hot path (default: ResNet-18, BASELINE configs[1]).

    python bench.py --gpus N --steps K --warmup W            (N>1: launched by torch.distributed.run)
    python bench.py --impl reference ...                      (the reference's own CPU path on the host cores)
    python bench.py --model {resnet18,vgg16,preact18,mobilenetv2}      (BASELINE configs 2-5)

One "step" = one pass of the hot path over one batch: augmentation -> forward -> loss -> backward ->
(gradient all-reduce) -> SGD -> weight re-layout. Synthetic 32x32x3 uint8 data, random-init weights (the package's
own seeded initialiser, engine.init_params). `value` is measured with 1024 images per GPU (weak scaling); with
N > 1 the same run also measures the reference's own partitioning — a global batch of 1024 divided across ranks,
main_dist.py:111 — and reports it under "strong". Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

UNIT = "images/s"
# SURVEY.md §8d algorithmic FLOPs per training image (2*M*N*K, unpadded K; fprop + dgrad + wgrad)
MODELS = {
    "resnet18": dict(spec="resnet_spec", flop=3.3290e9, batch=1024, bound="tensor", name="ResNet-18",
                     metric="resnet18_cifar10_train_images_per_sec", workload="resnet18_cifar10_train_batch1024"),
    "vgg16": dict(spec="vgg_spec", flop=1.8756e9, batch=1024, bound="tensor", name="VGG-16",
                  metric="vgg16_cifar10_train_images_per_sec", workload="vgg16_cifar10_train_batch1024"),
    "preact18": dict(spec="preact_resnet_spec", flop=3.3290e9, batch=1024, bound="tensor", name="PreActResNet-18",
                     metric="preactresnet18_cifar10_train_images_per_sec",
                     workload="preactresnet18_cifar10_train_batch1024"),
    "mobilenetv2": dict(spec="mobilenetv2_spec", flop=545.1e6, batch=512, bound="hbm", name="MobileNetV2",
                        metric="mobilenetv2_cifar10_train_images_per_sec",
                        workload="mobilenetv2_cifar10_train_batch512"),
}
NCU_TRAFFIC_FILE = os.path.join("profiles", "r2_ncu_conv_kernels.json")     # tools/ncu_summary.py on the GPU box


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return dict(hbm=p["hbm_gbs"], tf_burst=p["bf16_tflops"], tf_sustained=p["bf16_tflops_sustained"],
                    source="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            pass

    def stop(self):
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
            out, _ = self.proc.communicate()
        sm, mx, reasons = [], [], set()
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return dict(sm_mhz=sm[len(sm) // 2] if sm else None, sm_max_mhz=max(mx) if mx else None,
                    reasons=sorted(reasons), samples=len(sm))


# ----------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the reference's own CPU path, main.py:99-105 on models/resnet.py
# ----------------------------------------------------------------------------------------------
def cpu_reference_rate(steps, warmup, batch=128, pipeline_samples=0):
    """images/s of the reference's training-loop body (main.py:101-107) on the host cores, fp32, batch 128
    (BASELINE config 1). Uses the UNMODIFIED models/resnet.py when the reference files are on this machine
    (/root/reference, or the copy staged under oracle/_ref by oracle/stage_reference.py): kind "reference";
    otherwise the functional restatement oracle/nets.py: kind "port". pipeline_samples > 0 also times the
    reference's per-sample input pipeline (transform_train, main.py:30-35) on one core."""
    import numpy as np
    import torch
    from oracle.stage_reference import reference_dir
    torch.set_num_threads(os.cpu_count() or 1)
    rs = np.random.RandomState(0)
    x = torch.from_numpy(rs.standard_normal((batch, 3, 32, 32)).astype(np.float32))
    t = torch.from_numpy(rs.randint(0, 10, size=(batch,)).astype(np.int64))
    ref = reference_dir()
    if ref is not None:
        if ref not in sys.path:
            sys.path.insert(0, ref)
        import models.resnet as ref_resnet                                      # the reference's own file
        torch.manual_seed(0)
        net = ref_resnet.ResNet18()
        net.train()
        criterion = torch.nn.CrossEntropyLoss()                                 # main.py:86
        opt = torch.optim.SGD(net.parameters(), lr=0.1, momentum=0.9, weight_decay=5e-4)   # main.py:87-88
        forward, kind = (lambda: criterion(net(x), t)), "reference"
    else:
        from oracle import nets
        params, buffers = nets.build_resnet_params(0)
        leaves = {k: v.detach().clone().requires_grad_(True) for k, v in params.items()}
        opt = torch.optim.SGD(list(leaves.values()), lr=0.1, momentum=0.9, weight_decay=5e-4)
        forward = lambda: torch.nn.functional.cross_entropy(nets.resnet_forward(x, leaves, buffers, True), t)
        kind = "port"
    times = []
    for s in range(warmup + steps):
        t0 = time.perf_counter()
        opt.zero_grad()                                                    # main.py:101
        loss = forward()                                                   # :102-103
        loss.backward()                                                    # :104
        opt.step()                                                         # :105
        _ = loss.item()                                                    # :107
        if s >= warmup:
            times.append(time.perf_counter() - t0)
    total = sum(times)
    pipe = None
    if pipeline_samples > 0:
        try:
            import torchvision.transforms as T
            from PIL import Image
            tf = T.Compose([T.RandomCrop(32, padding=4), T.RandomHorizontalFlip(), T.ToTensor(),
                            T.Normalize((0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010))])     # main.py:30-35
            imgs = [Image.fromarray(rs.randint(0, 256, size=(32, 32, 3)).astype(np.uint8)) for _ in range(64)]
            t0 = time.perf_counter()
            for i in range(pipeline_samples):
                tf(imgs[i % 64])
            pipe = pipeline_samples / (time.perf_counter() - t0)
        except Exception:
            pipe = None
    return dict(rate=batch * len(times) / total, per_step=total / len(times), cores=torch.get_num_threads(),
                kind=kind, pipeline_samples_per_s_per_core=pipe)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    m = MODELS["resnet18"]
    steps, warm = min(args.steps, 10), min(args.warmup, 2)
    r = cpu_reference_rate(steps, warm, pipeline_samples=2000)
    what = "unmodified models/resnet.py ResNet18 + the loop body of main.py:101-107" if r["kind"] == "reference" \
        else "oracle/nets.py restatement of models/resnet.py + the loop body of main.py:101-107"
    sample = f"{steps} steps of batch 128 (one eighth of the batch-1024 workload step), fp32, {what}, synthetic tensors"
    line = {
        "impl": "reference", "metric": m["metric"], "value": r["rate"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warm, "ms_per_step": r["per_step"] * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": m["workload"], "model": m["name"], "sample_batch": 128},
        "cpu_baseline": {"value": r["rate"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"], "sample": sample,
                         "input_pipeline_samples_per_s_per_core": r["pipeline_samples_per_s_per_core"]},
        "e2e": {"value": r["rate"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------
# product arm
# ----------------------------------------------------------------------------------------------
LAUNCHES = {"b2c_conv_fprop": 1, "b2c_conv_wgrad": 1, "b2c_bn_finalize": 1, "b2c_bn_eval_scale_shift": 1,
            "b2c_bn_act_apply": 1, "b2c_bn_bwd_reduce": 1, "b2c_bn_bwd_finalize": 1, "b2c_bn_bwd_apply": 1,
            "b2c_head_fwd": 1, "b2c_ce_loss": 1, "b2c_head_bwd": 2, "b2c_augment_im2col": 1, "b2c_augment_nchw": 1,
            "b2c_stem_im2col": 1, "b2c_sgd_step": 1, "b2c_repack_weights": 1, "b2c_conv_dgrad_with_shortcut": 1,
            "b2c_conv_dgrad_bn": 1, "b2c_dwconv_fprop": 1, "b2c_dwconv_dgrad": 1, "b2c_dwconv_wgrad": 1,
            "b2c_maxpool2x2_fwd": 1, "b2c_maxpool2x2_bwd": 1, "b2c_bn_act_apply_stats": 1, "b2c_bn_bwd_apply_stats": 1,
            "b2c_zero": 1}


def count_launches(calls):
    n = 0
    for name, a in calls:
        if name == "b2c_conv_dgrad":
            n += 1          # (1x1 stride 2 without add_src also issues a cudaMemsetAsync of dx: not a kernel of ours)
        elif name == "b2c_conv_wgrad":
            d = a[0]._obj      # halo kernel + split reduction when a scratch buffer is passed and the shape qualifies
            halo = a[4] is not None and d.ksize == 3 and d.stride == 1 and d.h == d.w and d.w in (16, 32) \
                and ((d.cin == 64 and d.cout == 64) or (d.flags & 2)) and d.kflat_cin == 0 \
                and d.cin_valid in (0, d.cin) and d.cout_valid in (0, d.cout)
            n += 2 if halo else 1
        else:
            n += LAUNCHES.get(name, 1)
    return n


def conv_flops(calls):
    """Algorithmic FLOPs (2*M*N*K, unpadded K) of the recorded conv calls, split by kernel family."""
    tap = wg = 0.0
    for name, a in calls:
        if not name.startswith("b2c_conv_"):
            continue
        d = a[0]._obj
        ho, wo = d.h // d.stride, d.w // d.stride
        cin = d.cin_valid if d.cin_valid > 0 else d.cin
        fl = 2.0 * d.n * ho * wo * d.cout * cin * d.ksize * d.ksize
        if name == "b2c_conv_dgrad_with_shortcut":
            fl += 2.0 * d.n * ho * wo * d.cout * cin     # the fused 1x1 stride-2 shortcut dgrad
        if name == "b2c_conv_wgrad":
            wg += fl
        else:
            if name == "b2c_conv_fprop" and d.ksize == 1 and d.cin == 64 and d.h == 32 and d.stride == 1:
                fl = 2.0 * d.n * ho * wo * d.cout * 27      # im2col'ed stem: real K is 27
            tap += fl
    return tap, wg


def algorithmic_bytes_per_image(spec, engine):
    """HBM bytes per training image if every conv read its input once and wrote its output once, bf16 (SURVEY.md
    §8d 'fused-minimum conv traffic'), for forward, data gradient and weight gradient alike: 3 x sum over convs of
    (cin*H*W + cout*Ho*Wo) * 2 B, TRUE channel counts. The HBM-roofline numerator for MobileNetV2."""
    total = 0

    def conv(cin, cout, h, stride):
        nonlocal total
        total += (cin * h * h + cout * (h // stride) ** 2) * 2
    fam = spec["family"]
    if fam == "mobilenetv2":
        conv(3, 32, 32, 1)
        h = 32
        for b in spec["blocks"]:
            conv(b["cin"], b["planes"], h, 1)
            conv(b["planes"], b["planes"], h, b["stride"])
            h //= b["stride"]
            conv(b["planes"], b["cout"], h, 1)
            if b["shortcut"]:
                conv(b["cin"], b["cout"], h, 1)
        conv(320, 1280, h, 1)
    elif fam in ("resnet", "preact"):
        conv(3, 64, 32, 1)
        h = 32
        for b in spec["blocks"]:
            conv(b["cin"], b["cout"], h, b["stride"])
            if b["shortcut"]:
                conv(b["cin"], b["cout"], h, b["stride"])
            h //= b["stride"]
            conv(b["cout"], b["cout"], h, 1)
    elif fam == "vgg":
        h = 32
        for layer in spec["layers"]:
            if layer[0] == "pool":
                h //= 2
            else:
                conv(layer[3], layer[4], h, 1)
    return 3 * total


def run_product(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import b200cifar
    from b200cifar import _lib, data, engine, trainer

    M = MODELS[args.model]
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    batch = args.batch or M["batch"]
    K, W = args.steps, max(args.warmup, 3)
    spec = getattr(engine, M["spec"])()
    params, bufs = engine.init_params(spec, 0)             # random-init weights of the reference architecture
    n_data = 50000
    images, labels = data.synthetic_cifar(n_data, seed=0)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    def make_inputs(per_gpu):
        # this rank's sample order and crop/flip draws (DistributedSampler semantics when world > 1)
        idx_all = data.distributed_indices(n_data, world, rank) if world > 1 else \
            data.randperm(data.TorchCpuMT19937(0), n_data)
        rng = data.TorchCpuMT19937(1000 + rank)
        S = min(K + W, 64)                                  # distinct pre-staged steps (cycled)
        need = S * per_gpu
        idx_all = np.tile(idx_all, -(-need // len(idx_all)))[:need].reshape(S, per_gpu)
        ijf_all = data.draw_crop_flip(rng, need).reshape(S, per_gpu, 3)
        return S, idx_all, ijf_all

    def make_trainer(per_gpu, resident=True):
        tr = trainer.Trainer(spec, per_gpu, dev, world=world, rank=rank, use_graph=not args.no_graph,
                             bucket_bytes=int(args.bucket_mb * 2 ** 20), resident_dataset=resident)
        tr.load_params(params, bufs)
        if world > 1:
            dist.broadcast(tr.eng.P, 0)                     # DDP constructor semantics (main_dist.py:141)
            dist.broadcast(tr.eng.B, 0)
            tr.eng.repack()
        return tr

    def timed_resident(per_gpu, want_calls):
        """Device-resident arm at `per_gpu` images per rank: returns (ms per step, trainer, recorded calls, ...)."""
        S, idx_all, ijf_all = make_inputs(per_gpu)
        tr = make_trainer(per_gpu)
        tr.set_dataset(images, labels)
        d_idx = torch.from_numpy(idx_all).to(dev)
        d_ijf = torch.from_numpy(ijf_all).to(dev)
        d_tgt = torch.from_numpy(labels[idx_all]).to(dev)

        def resident_step(k):
            k %= S
            tr.idx_dev.copy_(d_idx[k]); tr.ijf_dev.copy_(d_ijf[k]); tr.target.copy_(d_tgt[k])
            tr._run()

        grad_ok = None
        if world > 1:
            # DDP's contract (main_dist.py:141-144): after backward every rank holds the MEAN of the per-rank
            # gradients. Checked on this very configuration before anything is timed: local gradients without the
            # all-reduce, gathered and averaged in fp64, against the arena after the bucketed NCCL path (which holds
            # the sum; the 1/world is applied inside the SGD kernel). BN buffers are restored afterwards.
            keepB = tr.eng.B.clone()
            tr.idx_dev.copy_(d_idx[0]); tr.ijf_dev.copy_(d_ijf[0]); tr.target.copy_(d_tgt[0])
            tr._forward_loss()
            tr.eng.backward()
            local_g = tr.eng.G.clone()
            allg = [torch.empty_like(local_g) for _ in range(world)]
            dist.all_gather(allg, local_g)
            mean = torch.stack([g.double() for g in allg]).mean(0)
            tr._forward_loss()
            tr._backward_reduce()
            torch.cuda.synchronize()
            err = ((tr.eng.G.double() / world - mean).norm() / mean.norm().clamp_min(1e-30)).item()
            distinct = (allg[0] - allg[-1]).abs().max().item() > 0        # the ranks really saw different samples
            flag = torch.tensor([1.0 if (err < 1e-4 and distinct) else 0.0], device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            grad_ok = dict(ok=bool(flag.item() == 1.0), rel_err=err, ranks_differ=bool(distinct))
            tr.eng.B.copy_(keepB)
            del allg, mean, local_g

        # first step eager with call recording: launch count + per-family replay lists
        _lib.recording = []
        use_graph = tr.use_graph
        tr.use_graph = False
        resident_step(0)
        calls = _lib.recording
        _lib.recording = None
        tr.use_graph = use_graph
        sampler = ClockSampler(local) if rank == 0 else None      # samples warm-up + timed region (all under load)
        for k in range(1, W):
            resident_step(k)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for k in range(W, W + K):
            resident_step(k)
        e1.record()
        barrier()
        ms = max_over_ranks(e0.elapsed_time(e1))
        clocks = sampler.stop() if sampler else None
        final_loss = tr.read_loss()
        in_sync = None
        if world > 1:      # DDP invariant: every replica holds the same parameters after training
            chk = torch.stack([tr.eng.P.double().sum(), tr.eng.P.double().abs().sum()])
            allchk = [torch.zeros_like(chk) for _ in range(world)]
            dist.all_gather(allchk, chk)
            in_sync = bool(all(torch.equal(allchk[0], c) for c in allchk))
        return dict(ms=ms, tr=tr, calls=calls if want_calls else None, clocks=clocks, final_loss=final_loss,
                    in_sync=in_sync, grad_ok=grad_ok, launches=count_launches(calls), inputs=(S, idx_all, ijf_all))

    # ---------------- device-resident arm: `value` (weak scaling: `batch` images per GPU)
    per_gpu = batch
    global_batch = per_gpu * world
    res = timed_resident(per_gpu, want_calls=True)
    ms, tr, calls, clocks = res["ms"], res["tr"], res["calls"], res["clocks"]
    value = global_batch * K / (ms * 1e-3)
    launches_per_step = res["launches"]
    S, idx_all, ijf_all = res["inputs"]

    # ---------------- roofline of the dominant kernel, timed live with CUDA events on the launching stream: all its
    # launches of one step, back to back, same buffers. The other conv kernel families are timed the same way.
    peaks = load_peaks()
    raw = _lib.load()
    tf_s, tf_b = peaks["tf_sustained"], peaks["tf_burst"]

    def is_halo(n, a):        # api.cu use_halo(): 3x3 stride-1 convs on 32/16/8-pixel images -> conv_halo2_kernel
        d = a[0]._obj
        return n in ("b2c_conv_fprop", "b2c_conv_dgrad", "b2c_conv_dgrad_bn") and d.ksize == 3 and d.stride == 1 \
            and d.h == d.w and d.w in (32, 16, 8) and not (d.flags & 1)

    def time_family(fam, reps=5):
        # the recorded argument tuples end with the stream the call was issued on (weight gradients: the engine's
        # side stream); for this measurement every launch goes to the bench stream, where the events are recorded
        import ctypes
        if not fam:
            return 0.0
        here = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        fam = [(n, tuple(a[:-1]) + (here,)) for n, a in fam]
        for n, a in fam:
            raw.raw(n)(*a)
        torch.cuda.synchronize()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record()
        for _ in range(reps):
            for n, a in fam:
                raw.raw(n)(*a)
        a1.record()
        torch.cuda.synchronize()
        return a0.elapsed_time(a1) / reps

    fd = ("b2c_conv_fprop", "b2c_conv_dgrad", "b2c_conv_dgrad_with_shortcut", "b2c_conv_dgrad_bn")
    halo_calls = [(n, a) for n, a in calls if n in fd and is_halo(n, a)]
    tap_calls = [(n, a) for n, a in calls if n in fd and not is_halo(n, a)]
    wg_calls = [(n, a) for n, a in calls if n == "b2c_conv_wgrad"]
    halo_fl, _ = conv_flops(halo_calls)
    tap_fl, _ = conv_flops(tap_calls)
    _, wg_fl = conv_flops(wg_calls)
    halo_ms, tap_ms, wg_ms = time_family(halo_calls), time_family(tap_calls), time_family(wg_calls)
    step_tf = value / world * M["flop"] / 1e12          # whole-step algorithmic TFLOP/s per GPU

    def fam(fl, ms_, n):
        a = fl / (ms_ * 1e-3) / 1e12 if ms_ > 0 else 0.0
        return {"achieved": a, "ms_per_step": ms_, "frac": a / tf_s, "frac_of_burst": a / tf_b, "launches_per_step": n}

    if M["bound"] == "tensor" and halo_calls:
        halo_launches = count_launches(halo_calls)
        achieved = halo_fl / (halo_ms * 1e-3) / 1e12
        traffic, traffic_src, tensor_pipe_pct = None, None, None
        try:        # DRAM bytes per launch of the same kernel from the committed ncu --set full capture
            with open(os.path.join(ROOT, NCU_TRAFFIC_FILE)) as f:
                cap = json.load(f)["conv_halo2_kernel"]
            traffic = cap["avg_dram_bytes_per_launch"]
            tensor_pipe_pct = cap.get("tensor_pipe_pct")
            traffic_src = f"{NCU_TRAFFIC_FILE} (dram__bytes_read.sum + dram__bytes_write.sum, average over the " \
                          f"{cap['launches']} conv_halo2_kernel launches of one batch-{cap.get('batch', 1024)} step)"
        except Exception:
            pass
        roofline = {"bound": "tensor",
                    "kernel": "conv_halo2_kernel (3x3 stride-1 conv fprop + dgrad, halo reuse, CTA pair / "
                              "tcgen05.mma.cta_group::2)",
                    "achieved": achieved, "peak": tf_s, "unit": "TFLOP/s", "frac": achieved / tf_s,
                    "peak_source": peaks["source"] + " bf16_tflops_sustained: the kernel is timed inside a loop of "
                                   "back-to-back launches (5 x its launches of one step, >= 7 ms), the sustained "
                                   "condition; against the burst figure the fraction is frac_of_burst",
                    "peak_burst": tf_b, "frac_of_burst": achieved / tf_b,
                    "traffic": traffic, "traffic_source": traffic_src,
                    "ncu_tensor_pipe_pct": tensor_pipe_pct,     # sm__pipe_tensor_cycles_active % of elapsed, same capture
                    "launches_per_step": halo_launches, "avg_launch_us": halo_ms * 1e3 / halo_launches,
                    "flop_per_launch_avg": halo_fl / halo_launches, "ms_per_step": halo_ms,
                    "tap_gemm": fam(tap_fl, tap_ms, count_launches(tap_calls)),
                    "wgrad": fam(wg_fl, wg_ms, count_launches(wg_calls)),
                    "conv_ms_per_step": halo_ms + tap_ms + wg_ms}
    elif M["bound"] == "tensor":
        roofline = {"bound": "tensor", "kernel": "whole step (all conv kernels)", "achieved": step_tf, "peak": tf_s,
                    "unit": "TFLOP/s", "frac": step_tf / tf_s, "peak_burst": tf_b, "frac_of_burst": step_tf / tf_b,
                    "traffic": None, "tap_gemm": fam(tap_fl, tap_ms, count_launches(tap_calls)),
                    "wgrad": fam(wg_fl, wg_ms, count_launches(wg_calls))}
    else:
        bpi = algorithmic_bytes_per_image(spec, engine)
        gbs = value / world * bpi / 1e9
        roofline = {"bound": "hbm", "kernel": "whole step (depthwise + pointwise conv, BatchNorm passes)",
                    "achieved": gbs, "peak": peaks["hbm"], "unit": "GB/s", "frac": gbs / peaks["hbm"],
                    "peak_source": peaks["source"] + " hbm_gbs", "traffic": None,
                    "algorithmic_bytes_per_image": bpi,
                    "note": "algorithmic bytes = 3 x sum over convs of (input + output) bf16 bytes, true channel "
                            "counts (SURVEY.md §8d fused minimum); BatchNorm passes and channel padding are overhead",
                    "tensor_tflops_whole_step": step_tf}
    del tr, res["tr"]
    torch.cuda.empty_cache()

    # ---------------- the reference's own partitioning: global batch divided across ranks (main_dist.py:111)
    strong = None
    if world > 1 and not args.no_strong and batch % world == 0:
        sres = timed_resident(batch // world, want_calls=False)
        strong = {"value": batch * K / (sres["ms"] * 1e-3), "unit": UNIT, "ms_per_step": sres["ms"] / K,
                  "per_gpu_batch": batch // world, "global_batch": batch, "launches_per_step": sres["launches"],
                  "replicas_in_sync": sres["in_sync"], "grad_mean_ok": sres["grad_ok"],
                  "clocks": sres["clocks"], "final_loss": sres["final_loss"],
                  "note": "scaling 'strong': total work fixed at the batch-1024 step of main_dist.py; "
                          "efficiency = value / (n_gpus * the n_gpus=1 value)"}
        del sres
        torch.cuda.empty_cache()

    # ---------------- end-to-end arm: host uint8 batches through the public API, H2D + loss D2H every step
    tr2 = make_trainer(per_gpu, resident=False)
    nb = min(S, 8)
    host_batches = [np.ascontiguousarray(images[idx_all[k]]) for k in range(nb)]
    host_labels = [labels[idx_all[k]] for k in range(nb)]
    for k in range(W):
        tr2.step_host_batch(host_batches[k % nb], host_labels[k % nb], ijf_all[k % S])
        tr2.read_loss()
    # timed region: every step = swap-in of the staged batch + the step + staging of the NEXT host batch (pinned copy
    # and H2D on a copy stream, overlapping the step the way DataLoader workers do, main.py:44-45) + loss D2H
    j = W % nb
    tr2.prefetch_host_batch(host_batches[j], host_labels[j], ijf_all[W % S])
    barrier()
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    for k in range(W, W + K):
        tr2.step_prefetched()
        j = (k + 1) % nb
        tr2.prefetch_host_batch(host_batches[j], host_labels[j], ijf_all[(k + 1) % S])
        tr2.read_loss()
    t1.record()
    barrier()
    ms2 = max_over_ranks(t0.elapsed_time(t1))
    e2e = {"value": global_batch * K / (ms2 * 1e-3), "unit": UNIT,
           "h2d_bytes_per_step": per_gpu * (3072 + 3 + 8), "d2h_bytes_per_step": 4,
           "ms_per_step": ms2 / K}
    del tr2

    if rank == 0:
        cpu = cpu_reference_rate(3, 1, pipeline_samples=1000) if (world == 1 and not args.no_cpu) else None
        line = {
            "metric": M["metric"], "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None,      # BASELINE.md holds no published number for this metric (the reference has none)
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": M["workload"], "model": M["name"],
                       "per_gpu_batch": per_gpu, "global_batch": global_batch, "parallelism": f"dp{world}",
                       "l2": "working set per step (activations > 2 GB) exceeds the 126 MB L2; no flush needed",
                       "cuda_graph": not args.no_graph},
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches_per_step * K,
            "launches_per_step": launches_per_step, "roofline": roofline,
            "tensor_frac_whole_step": step_tf / tf_s, "tensor_frac_whole_step_of_burst": step_tf / tf_b,
            "final_loss": res["final_loss"], "replicas_in_sync": res["in_sync"],
        }
        if res["grad_ok"] is not None:
            line["grad_mean_ok"] = res["grad_ok"]["ok"]
            line["grad_mean_check"] = res["grad_ok"]
        if strong is not None:
            line["strong"] = strong
        if cpu is not None:
            line["cpu_baseline"] = {
                "value": cpu["rate"], "unit": UNIT, "cores": cpu["cores"], "kind": cpu["kind"],
                "sample": "3 steps of batch 128 (BASELINE config 1), fp32, " +
                          ("unmodified models/resnet.py" if cpu["kind"] == "reference" else "oracle/nets.py") +
                          " under the loop body of main.py:101-107",
                "input_pipeline_samples_per_s_per_core": cpu["pipeline_samples_per_s_per_core"]}
            line["gpu_over_cpu_baseline"] = value / cpu["rate"]
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default="resnet18", choices=sorted(MODELS))
    ap.add_argument("--batch", type=int, default=0, help="images per GPU (default: 1024; MobileNetV2 512)")
    ap.add_argument("--no-strong", action="store_true", help="N > 1: skip the global-batch-1024 (strong scaling) leg")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--bucket-mb", type=float, default=25.0, help="gradient all-reduce bucket size (MiB), N > 1")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_product(args)


if __name__ == "__main__":
    main()
