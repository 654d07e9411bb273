"""Whole-step trainer: the loop body of the reference's train() (main.py:99-105; main_dist.py:174-191)

    inputs, targets -> optimizer.zero_grad() -> net(inputs) -> criterion -> loss.backward() -> optimizer.step()

as one static sequence of kernel launches, optionally captured into a single CUDA graph. The input
pipeline (main.py:30-35) runs on the GPU over a uint8 dataset resident in HBM, driven by host-side draws
that reproduce torch's CPU generator (data.py). With world > 1 the gradient arena is all-reduced with
NCCL in buckets on a side stream while backward continues (the DDP Reducer's job, main_dist.py:141-144);
the 1/world of DDP's mean is folded into the SGD kernel.
"""
import numpy as np
import torch

from . import data as hostdata
from . import ops
from .engine import Engine, plan_grad_buckets


class Trainer:
    def __init__(self, spec, batch, device="cuda", lr=0.1, momentum=0.9, weight_decay=5e-4, world=1, rank=0,
                 use_graph=True, resident_dataset=True, bucket_bytes=25 * 2 ** 20):
        self.device = torch.device(device)
        self.N, self.world, self.rank = batch, world, rank
        self.eng = Engine(spec, batch, device)
        self.K = spec["num_classes"]
        dev = self.device
        self.hyper_host = torch.tensor([lr, momentum, weight_decay, 1.0 / world, 1.0], dtype=torch.float32).pin_memory()
        self.hyper = self.hyper_host.to(dev)
        self.idx_dev = torch.zeros(batch, dtype=torch.int64, device=dev)
        self.ijf_dev = torch.zeros(batch, 3, dtype=torch.int8, device=dev)
        self.target = torch.zeros(batch, dtype=torch.int64, device=dev)
        self.loss = torch.zeros(1, dtype=torch.float32, device=dev)
        self.correct = torch.zeros(1, dtype=torch.int32, device=dev)
        self.resident = resident_dataset
        self.data = None
        if not resident_dataset:
            # e2e mode: the caller hands over host batches; they are staged here and read with idx = 0..N-1
            self.data = torch.zeros(batch, 32, 32, 3, dtype=torch.uint8, device=dev)
            self.idx_dev.copy_(torch.arange(batch, device=dev))
        # pinned staging ring for per-step host inputs
        self._ring = 4
        self._slot = 0
        self._pin_idx = [torch.zeros(batch, dtype=torch.int64).pin_memory() for _ in range(self._ring)]
        self._pin_ijf = [torch.zeros(batch, 3, dtype=torch.int8).pin_memory() for _ in range(self._ring)]
        self._pin_tgt = [torch.zeros(batch, dtype=torch.int64).pin_memory() for _ in range(self._ring)]
        self._pin_img = None
        self._ring_events = [None] * self._ring
        self._loss_host = torch.zeros(1, dtype=torch.float32).pin_memory()
        self.use_graph = use_graph
        self.graph = None
        self.steps_done = 0
        self.launches_per_step = None
        self.comm_stream = torch.cuda.Stream(device=dev) if world > 1 else None
        self.bucket_bytes = bucket_bytes
        self._buckets = self._plan_buckets() if world > 1 else []

    # ------------------------------------------------------------------ setup
    def load_params(self, params, buffers=None):
        self.eng.load_params(params, buffers)
        if buffers is None:
            self.eng.init_bn_buffers()

    def set_dataset(self, images_u8, labels):
        """images_u8: [n,32,32,3] uint8 (numpy or tensor), copied to HBM once; labels stay on the host."""
        assert self.resident
        t = torch.as_tensor(images_u8)
        assert t.dtype == torch.uint8 and tuple(t.shape[1:]) == (32, 32, 3)
        self.data = t.contiguous().to(self.device)
        self.labels_host = np.asarray(labels, dtype=np.int64)

    def set_lr(self, lr):
        self.hyper_host[0] = lr
        self.hyper.copy_(self.hyper_host, non_blocking=True)

    # ------------------------------------------------------------------ gradient buckets (DDP)
    def _plan_buckets(self):
        buckets, ready = plan_grad_buckets(self.eng.spec, self.bucket_bytes)
        self._bucket_ready_after = ready
        return buckets

    # ------------------------------------------------------------------ one step
    def _launch_step(self):
        eng = self.eng
        ops.augment_im2col(self.data, self.idx_dev, self.ijf_dev, eng.x0)
        eng.zero_grad()
        eng.forward(True)
        ops.ce_loss(eng.logits, self.target, self.N, self.K, eng.dlogits, self.loss, self.correct)
        if self.world == 1:
            eng.backward()
        else:
            import torch.distributed as dist
            main = torch.cuda.current_stream()
            for pos, op in enumerate(eng.bwd_ops):
                op()
                bi = self._bucket_ready_after.get(pos)
                if bi is not None:
                    a, b = self._buckets[bi]
                    ev = torch.cuda.Event()
                    ev.record(main)
                    self.comm_stream.wait_event(ev)
                    side = eng.side_marker()              # weight gradients of this bucket run on the side stream
                    if side is not None:
                        self.comm_stream.wait_event(side)
                    with torch.cuda.stream(self.comm_stream):
                        dist.all_reduce(eng.G[a:b])      # sum; the mean's 1/world is applied inside sgd_step
            eng.join_side()
            done = torch.cuda.Event()
            done.record(self.comm_stream)
            main.wait_event(done)
        ops.sgd_step(eng.P, eng.G, eng.M, self.hyper)
        eng.repack()

    def _stage_inputs(self, idx, ijf, targets, images=None):
        s = self._slot
        self._slot = (s + 1) % self._ring
        if self._ring_events[s] is not None:
            self._ring_events[s].synchronize()
        n = self.N
        self._pin_ijf[s].numpy()[...] = ijf
        self._pin_tgt[s].numpy()[...] = targets
        self.ijf_dev.copy_(self._pin_ijf[s], non_blocking=True)
        self.target.copy_(self._pin_tgt[s], non_blocking=True)
        if images is not None:
            if self._pin_img is None:
                self._pin_img = [torch.zeros(n, 32, 32, 3, dtype=torch.uint8).pin_memory() for _ in range(self._ring)]
            self._pin_img[s].numpy()[...] = images
            self.data.copy_(self._pin_img[s], non_blocking=True)
        else:
            self._pin_idx[s].numpy()[...] = idx
            self.idx_dev.copy_(self._pin_idx[s], non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self._ring_events[s] = ev

    def _run(self):
        if not self.use_graph:
            self._launch_step()
        else:
            if self.graph is None:
                self._capture()
            self.graph.replay()
        if self.steps_done == 0:
            self.hyper_host[4] = 0.0
            self.hyper.copy_(self.hyper_host, non_blocking=True)
        self.steps_done += 1

    def _capture(self):
        # warm-up outside capture is the caller's first eager step(s); here: capture only.
        # State mutated by a step (params, momentum, BN stats) is advanced by replays only, so save/restore
        # is unnecessary: capture itself does not execute kernels.
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._launch_step()
        self.graph = g

    def step(self, idx, ijf):
        """One training step on dataset samples `idx` (host int64 [N]) with crop/flip draws `ijf` (int8 [N,3])."""
        self._stage_inputs(idx, ijf, self.labels_host[idx])
        self._run()
        return self.loss

    def step_host_batch(self, images_u8, labels, ijf):
        """One training step on a HOST uint8 batch [N,32,32,3] (copied to the device inside the call)."""
        assert not self.resident
        self._stage_inputs(None, ijf, labels, images=images_u8)
        self._run()
        return self.loss

    # ------------------------------------------------------------------ pipelined host input (DataLoader prefetch)
    def prefetch_host_batch(self, images_u8, labels, ijf):
        """Stage the NEXT step's host batch while the current step runs on the GPU — what the reference gets from
        DataLoader(num_workers=2) + `.to(device)` (main.py:44-45,100): host copy into pinned memory, then H2D on a
        copy stream into staging buffers. `step_prefetched()` swaps them in (device-to-device) and runs the step."""
        assert not self.resident
        if not hasattr(self, "_copy_stream"):
            dev = self.device
            self._copy_stream = torch.cuda.Stream(device=dev)
            self._stg_img = torch.zeros_like(self.data)
            self._stg_ijf = torch.zeros_like(self.ijf_dev)
            self._stg_tgt = torch.zeros_like(self.target)
            self._h2d_done, self._swap_done = torch.cuda.Event(), None
        s = self._slot
        self._slot = (s + 1) % self._ring
        if self._ring_events[s] is not None:
            self._ring_events[s].synchronize()
        if self._pin_img is None:
            self._pin_img = [torch.zeros(self.N, 32, 32, 3, dtype=torch.uint8).pin_memory() for _ in range(self._ring)]
        self._pin_img[s].numpy()[...] = images_u8
        self._pin_ijf[s].numpy()[...] = ijf
        self._pin_tgt[s].numpy()[...] = labels
        cs = self._copy_stream
        if self._swap_done is not None:
            cs.wait_event(self._swap_done)            # the previous swap has read the staging buffers
        with torch.cuda.stream(cs):
            self._stg_img.copy_(self._pin_img[s], non_blocking=True)
            self._stg_ijf.copy_(self._pin_ijf[s], non_blocking=True)
            self._stg_tgt.copy_(self._pin_tgt[s], non_blocking=True)
            self._h2d_done.record(cs)
            ev = torch.cuda.Event()
            ev.record(cs)
        self._ring_events[s] = ev

    def step_prefetched(self):
        """One training step on the batch staged by the last prefetch_host_batch()."""
        main = torch.cuda.current_stream()
        main.wait_event(self._h2d_done)
        self.data.copy_(self._stg_img, non_blocking=True)
        self.ijf_dev.copy_(self._stg_ijf, non_blocking=True)
        self.target.copy_(self._stg_tgt, non_blocking=True)
        self._swap_done = torch.cuda.Event()
        self._swap_done.record(main)
        self._run()
        return self.loss

    # ------------------------------------------------------------------ evaluation (test(), main.py:116-134)
    def evaluate(self, images_u8, labels, batch=100):
        """Top-1 accuracy and mean loss over a host uint8 set with transform_test (ToTensor + Normalize only,
        main.py:37-40), BatchNorm in eval mode (running statistics), batches of `batch` like the reference's
        testloader (main.py:49-50). Runs on a second plan that shares this trainer's parameter arenas."""
        images_u8 = np.ascontiguousarray(images_u8)
        labels = np.asarray(labels, dtype=np.int64)
        n = len(labels)
        correct, loss_sum = 0, 0.0
        for s in range(0, n, batch):
            b = min(batch, n - s)
            ev = self._eval_plans.get(b) if hasattr(self, "_eval_plans") else None
            if ev is None:
                if not hasattr(self, "_eval_plans"):
                    self._eval_plans = {}
                eng = Engine(self.eng.spec, b, self.device, share_from=self.eng)
                ev = dict(eng=eng, data=torch.zeros(b, 32, 32, 3, dtype=torch.uint8, device=self.device),
                          idx=torch.arange(b, device=self.device),
                          ijf=torch.tensor([[4, 4, 0]] * b, dtype=torch.int8, device=self.device),  # identity crop
                          tgt=torch.zeros(b, dtype=torch.int64, device=self.device),
                          loss=torch.zeros(1, device=self.device),
                          correct=torch.zeros(1, dtype=torch.int32, device=self.device))
                self._eval_plans[b] = ev
            eng = ev["eng"]
            ev["data"].copy_(torch.from_numpy(images_u8[s:s + b]))
            ev["tgt"].copy_(torch.from_numpy(labels[s:s + b]))
            ops.augment_im2col(ev["data"], ev["idx"], ev["ijf"], eng.x0)
            eng.forward(False)
            ops.ce_loss(eng.logits, ev["tgt"], b, self.K, eng.dlogits, ev["loss"], ev["correct"])
            correct += int(ev["correct"].item())
            loss_sum += float(ev["loss"].item()) * b
        return correct / n, loss_sum / n

    def read_loss(self):
        """Device -> host read of the last step's loss (the reference's loss.item(), main.py:107)."""
        self._loss_host.copy_(self.loss, non_blocking=False)
        return float(self._loss_host[0])


def make_epoch_plan(n, batch, world=1, rank=0, seed=0, distributed=False):
    """Index batches for one epoch the way the reference's loaders produce them."""
    if distributed:
        idx = hostdata.distributed_indices(n, world, rank, seed=0, epoch=0)
    else:
        idx = hostdata.randperm(hostdata.TorchCpuMT19937(seed), n)
    return hostdata.batches(idx, batch)
