#!/usr/bin/env python
"""bench.py — ResNet-18 CIFAR-10 training throughput of the B200-native hot path.

    python bench.py --gpus N --steps K --warmup W            (N>1: launched by torch.distributed.run)
    python bench.py --impl reference ...                      (the reference's CPU path, oracle port)

One "step" = one pass of the hot path over one batch: augmentation -> forward -> loss -> backward ->
(gradient all-reduce) -> SGD -> weight re-layout. Workload: BASELINE.json configs[1], ResNet-18, batch 1024
per GPU (weak scaling; `--scaling strong` divides a global batch of 1024 across ranks the way
main_dist.py:111 does), synthetic 32x32x3 uint8 data, random-init weights.
Prints ONE JSON line on rank 0.
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

RESNET18_TRAIN_FLOP_PER_IMG = 3.3290e9      # fprop 1110.8 + dgrad 1107.3 + wgrad 1110.8 MFLOP
METRIC = "resnet18_cifar10_train_images_per_sec"
UNIT = "images/s"


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
# reference arm / cpu baseline: the reference's CPU path (oracle port of main.py:99-105 + models/resnet.py)
# ----------------------------------------------------------------------------------------------
def cpu_reference_rate(steps, warmup, batch=128):
    import numpy as np
    import torch
    from oracle import nets
    torch.set_num_threads(os.cpu_count() or 1)
    params, buffers = nets.build_resnet_params(0)
    leaves = {k: v.detach().clone().requires_grad_(True) for k, v in params.items()}
    opt = torch.optim.SGD(list(leaves.values()), lr=0.1, momentum=0.9, weight_decay=5e-4)   # main.py:87-88
    rs = np.random.RandomState(0)
    x = torch.from_numpy(rs.standard_normal((batch, 3, 32, 32)).astype(np.float32))
    t = torch.from_numpy(rs.randint(0, 10, size=(batch,)).astype(np.int64))
    times = []
    for s in range(warmup + steps):
        t0 = time.perf_counter()
        opt.zero_grad()                                                    # main.py:101
        loss = torch.nn.functional.cross_entropy(nets.resnet_forward(x, leaves, buffers, True), t)   # :102-103
        loss.backward()                                                    # :104
        opt.step()                                                         # :105
        _ = loss.item()                                                    # :107
        if s >= warmup:
            times.append(time.perf_counter() - t0)
    total = sum(times)
    return batch * len(times) / total, total / len(times), torch.get_num_threads()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = min(args.steps, 10)
    rate, per_step, cores = cpu_reference_rate(steps, min(args.warmup, 2))
    sample = f"{steps} steps of batch 128 (one eighth of the batch-1024 workload step), fp32, torch CPU ops"
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": min(args.warmup, 2), "ms_per_step": per_step * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "resnet18_cifar10_train_batch1024", "model": "ResNet-18", "sample_batch": 128},
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------
# product arm
# ----------------------------------------------------------------------------------------------
LAUNCHES = {"b2c_conv_fprop": 1, "b2c_conv_wgrad": 1, "b2c_bn_finalize": 1, "b2c_bn_eval_scale_shift": 1,
            "b2c_bn_act_apply": 1, "b2c_bn_bwd_reduce": 1, "b2c_bn_bwd_finalize": 1, "b2c_bn_bwd_apply": 1,
            "b2c_head_fwd": 1, "b2c_ce_loss": 1, "b2c_head_bwd": 2, "b2c_augment_im2col": 1, "b2c_augment_nchw": 1,
            "b2c_stem_im2col": 1, "b2c_sgd_step": 1, "b2c_repack_weights": 1, "b2c_conv_dgrad_with_shortcut": 1, "b2c_conv_dgrad_bn": 1}


def count_launches(calls):
    n = 0
    for name, a in calls:
        if name == "b2c_conv_dgrad":
            d = a[0]._obj
            # 1x1 stride 2 without add_src: cudaMemsetAsync of dx (not a kernel of ours) + one launch
            n += 1
        elif name == "b2c_conv_wgrad":
            d = a[0]._obj      # halo kernel + split reduction when a scratch buffer is passed and the shape qualifies
            halo = a[4] is not None and d.ksize == 3 and d.stride == 1 and d.h == d.w and d.w in (16, 32) \
                and ((d.cin == 64 and d.cout == 64) or (d.flags & 2)) and d.kflat_cin == 0 and d.cin_valid in (0, d.cin) and d.cout_valid in (0, d.cout)
            n += 2 if halo else 1
        else:
            n += LAUNCHES[name]
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


def run_product(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import b200cifar
    from b200cifar import _lib, data, engine, trainer
    from oracle import nets

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    per_gpu = args.batch if args.scaling == "weak" else args.batch // world
    global_batch = per_gpu * world
    K, W = args.steps, max(args.warmup, 3)

    params, buffers = nets.build_resnet_params(0)            # random-init weights of the reference architecture
    bufs = {k: v for k, v in buffers.items() if not k.endswith("num_batches_tracked")}
    n_data = 50000
    images, labels = data.synthetic_cifar(n_data, seed=0)
    # this rank's sample order and crop/flip draws (DistributedSampler semantics when world > 1)
    idx_all = data.distributed_indices(n_data, world, rank) if world > 1 else data.randperm(data.TorchCpuMT19937(0), n_data)
    rng = data.TorchCpuMT19937(1000 + rank)
    S = min(K + W, 64)                                  # distinct pre-staged steps (cycled)
    need = S * per_gpu
    reps = -(-need // len(idx_all))
    idx_all = np.tile(idx_all, reps)[:need].reshape(S, per_gpu)
    ijf_all = data.draw_crop_flip(rng, need).reshape(S, per_gpu, 3)

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

    # ---------------- device-resident arm: `value`
    tr = trainer.Trainer(engine.resnet_spec(), per_gpu, dev, world=world, rank=rank, use_graph=not args.no_graph,
                         bucket_bytes=int(args.bucket_mb * 2 ** 20))
    tr.load_params(params, bufs)
    if world > 1:
        dist.broadcast(tr.eng.P, 0)
        dist.broadcast(tr.eng.B, 0)
        tr.eng.repack()
    tr.set_dataset(images, labels)
    d_idx = torch.from_numpy(idx_all).to(dev)
    d_ijf = torch.from_numpy(ijf_all).to(dev)
    d_tgt = torch.from_numpy(labels[idx_all]).to(dev)

    def resident_step(k):
        k %= S
        tr.idx_dev.copy_(d_idx[k]); tr.ijf_dev.copy_(d_ijf[k]); tr.target.copy_(d_tgt[k])
        tr._run()

    # first step eager with call recording: launch count + per-family replay lists
    _lib.recording = []
    use_graph = tr.use_graph
    tr.use_graph = False
    resident_step(0)
    calls = _lib.recording
    _lib.recording = None
    tr.use_graph = use_graph
    launches_per_step = count_launches(calls)
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
    value = global_batch * K / (ms * 1e-3)
    final_loss = tr.read_loss()
    in_sync = None
    if world > 1:      # DDP invariant (main_dist.py:141-144): every replica holds the same parameters after training
        chk = torch.stack([tr.eng.P.double().sum(), tr.eng.P.double().abs().sum()])
        allchk = [torch.zeros_like(chk) for _ in range(world)]
        dist.all_gather(allchk, chk)
        in_sync = bool(all(torch.equal(allchk[0], c) for c in allchk))

    # ---------------- roofline of the dominant kernel (conv_halo_kernel: every 3x3 stride-1 fprop / dgrad), timed
    # live with CUDA events on the launching stream: all its launches of one step, back to back, same buffers.
    # The other two conv kernel families (per-tap GEMM, wgrad) are timed the same way and reported beside it.
    peaks = load_peaks()
    raw = _lib.load()

    def is_halo(n, a):        # api.cu use_halo(): 3x3 stride-1 convs on 32/16/8-pixel images -> conv_halo_kernel
        d = a[0]._obj
        return n in ("b2c_conv_fprop", "b2c_conv_dgrad", "b2c_conv_dgrad_bn") and d.ksize == 3 and d.stride == 1 \
            and d.h == d.w and d.w in (32, 16, 8) and not (d.flags & 1)

    fd = ("b2c_conv_fprop", "b2c_conv_dgrad", "b2c_conv_dgrad_with_shortcut", "b2c_conv_dgrad_bn")
    halo_calls = [(n, a) for n, a in calls if n in fd and is_halo(n, a)]
    tap_calls = [(n, a) for n, a in calls if n in fd and not is_halo(n, a)]
    wg_calls = [(n, a) for n, a in calls if n == "b2c_conv_wgrad"]
    halo_fl, _ = conv_flops(halo_calls)
    tap_fl, _ = conv_flops(tap_calls)
    _, wg_fl = conv_flops(wg_calls)

    def time_family(fam, reps=5):
        # the recorded argument tuples end with the stream the call was issued on (weight gradients: the engine's
        # side stream); for this measurement every launch goes to the bench stream, where the events are recorded
        import ctypes
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

    halo_ms = time_family(halo_calls)
    tap_ms = time_family(tap_calls)
    wg_ms = time_family(wg_calls)
    halo_launches = count_launches(halo_calls)
    achieved = halo_fl / (halo_ms * 1e-3) / 1e12
    traffic = None          # DRAM bytes per launch of the same kernel from the committed ncu --set full capture
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles",
                               "r1d_ncu_conv_halo_step.json")) as f:
            traffic = json.load(f)["avg_dram_bytes_per_launch"]
    except Exception:
        pass
    tf = peaks["tf_sustained"]
    roofline = {"bound": "tensor", "kernel": "conv_halo2_kernel (3x3 stride-1 conv fprop + dgrad, halo reuse, CTA pair / tcgen05.mma.cta_group::2)",
                "achieved": achieved, "peak": tf, "unit": "TFLOP/s", "frac": achieved / tf,
                "traffic": traffic, "traffic_source": "profiles/r1d_ncu_conv_halo_step.json (dram read+write, avg per launch; captured on the single-CTA conv_halo_kernel, same tensors and tiles)",
                "peak_source": peaks["source"] + " bf16_tflops_sustained",
                "launches_per_step": halo_launches, "avg_launch_us": halo_ms * 1e3 / halo_launches,
                "flop_per_launch_avg": halo_fl / halo_launches, "ms_per_step": halo_ms,
                "tap_gemm": {"achieved": tap_fl / (tap_ms * 1e-3) / 1e12, "ms_per_step": tap_ms,
                             "frac": tap_fl / (tap_ms * 1e-3) / 1e12 / tf,
                             "launches_per_step": count_launches(tap_calls)},
                "wgrad": {"achieved": wg_fl / (wg_ms * 1e-3) / 1e12, "ms_per_step": wg_ms,
                          "frac": wg_fl / (wg_ms * 1e-3) / 1e12 / tf, "launches_per_step": count_launches(wg_calls)},
                "conv_ms_per_step": halo_ms + tap_ms + wg_ms}
    del tr
    torch.cuda.empty_cache()

    # ---------------- end-to-end arm: host uint8 batches through the public API, H2D + loss D2H every step
    tr2 = trainer.Trainer(engine.resnet_spec(), per_gpu, dev, world=world, rank=rank, use_graph=not args.no_graph,
                          bucket_bytes=int(args.bucket_mb * 2 ** 20),
                          resident_dataset=False)
    tr2.load_params(params, bufs)
    host_batches = [np.ascontiguousarray(images[idx_all[k]]) for k in range(min(S, 8))]
    host_labels = [labels[idx_all[k]] for k in range(min(S, 8))]
    for k in range(W):
        tr2.step_host_batch(host_batches[k % len(host_batches)], host_labels[k % len(host_batches)], ijf_all[k % S])
        tr2.read_loss()
    # timed region: every step = swap-in of the staged batch + the step + staging of the NEXT host batch (pinned copy
    # and H2D on a copy stream, overlapping the step the way DataLoader workers do, main.py:44-45) + loss D2H
    j = W % len(host_batches)
    tr2.prefetch_host_batch(host_batches[j], host_labels[j], ijf_all[W % S])
    barrier()
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    for k in range(W, W + K):
        tr2.step_prefetched()
        j = (k + 1) % len(host_batches)
        tr2.prefetch_host_batch(host_batches[j], host_labels[j], ijf_all[(k + 1) % S])
        tr2.read_loss()
    t1.record()
    barrier()
    ms2 = max_over_ranks(t0.elapsed_time(t1))
    e2e = {"value": global_batch * K / (ms2 * 1e-3), "unit": UNIT,
           "h2d_bytes_per_step": per_gpu * (3072 + 3 + 8), "d2h_bytes_per_step": 4,
           "ms_per_step": ms2 / K}
    del tr2

    line = None
    if rank == 0:
        cpu_rate, cpu_step, cores = cpu_reference_rate(3, 1) if (world == 1 and not args.no_cpu) else (None, None, None)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "resnet18_cifar10_train_batch1024", "model": "ResNet-18",
                       "per_gpu_batch": per_gpu, "global_batch": global_batch, "parallelism": f"dp{world}",
                       "l2": "working set per step (activations > 2 GB) exceeds the 126 MB L2; no flush needed",
                       "cuda_graph": not args.no_graph},
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches_per_step * K,
            "launches_per_step": launches_per_step, "roofline": roofline,
            "tensor_frac_whole_step": value / world * RESNET18_TRAIN_FLOP_PER_IMG / 1e12 / peaks["tf_sustained"],
            "final_loss": final_loss, "replicas_in_sync": in_sync,
        }
        if cpu_rate is not None:
            line["cpu_baseline"] = {"value": cpu_rate, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": "3 steps of batch 128 (BASELINE config 1), fp32 torch CPU ops via "
                                              "oracle/nets.py replaying main.py:99-105"}
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
    ap.add_argument("--batch", type=int, default=1024, help="per-GPU batch (weak) or global batch (strong)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
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
