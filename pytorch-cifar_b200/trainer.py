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
        # {lr, momentum, weight_decay, 1/world, first_step}: host values in pageable memory; every update is a
        # blocking 20-byte copy of a private snapshot, so a value written here can never reach steps that were
        # enqueued earlier (an async copy from one shared pinned buffer could)
        self.hyper_host = torch.tensor([lr, momentum, weight_decay, 1.0 / world, 1.0], dtype=torch.float32)
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
        # launch plans by batch size: the full batch, plus one per ragged last batch met so far (the reference's
        # loaders keep it: drop_last=False, 50000 = 390*128 + 80, main.py:45; 6250 = 48*128 + 106 per rank,
        # main_dist.py:111). A tail plan shares every parameter / gradient / momentum / BN-buffer arena.
        self._main = dict(eng=self.eng, idx=self.idx_dev, ijf=self.ijf_dev, tgt=self.target, graph=None)
        self._plans = {batch: self._main}
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

    def _push_hyper(self):
        self.hyper.copy_(self.hyper_host.clone(), non_blocking=False)

    def set_lr(self, lr):
        """Learning rate for the steps issued from now on (the scheduler stays on the host, main.py:89,154)."""
        self.hyper_host[0] = lr
        self._push_hyper()

    # ------------------------------------------------------------------ gradient buckets (DDP)
    def _plan_buckets(self):
        buckets, ready = plan_grad_buckets(self.eng.spec, self.bucket_bytes)
        self._bucket_ready_after = ready
        return buckets

    # ------------------------------------------------------------------ one step
    def _forward_loss(self, pl=None):
        pl = pl or self._main
        eng = pl["eng"]
        ops.augment_im2col(self.data, pl["idx"], pl["ijf"], eng.x0, taps=eng.stem_taps)
        eng.zero_grad()
        eng.forward(True)
        ops.ce_loss(eng.logits, pl["tgt"], eng.N, self.K, eng.dlogits, self.loss, self.correct)

    def _backward_reduce(self, pl=None):
        """loss.backward() including DDP's gradient all-reduce; leaves the SUM over ranks in the arena."""
        eng = (pl or self._main)["eng"]
        if self.world == 1:
            eng.backward()
        else:
            eng.backward_allreduce(self._buckets, self._bucket_ready_after, self.comm_stream)

    def _optimizer_step(self, pl=None):
        eng = (pl or self._main)["eng"]
        ops.sgd_step(eng.P, eng.G, eng.M, self.hyper)      # hyper[3] = 1/world: DDP's mean
        eng.repack()

    def _launch_step(self, pl=None):
        self._forward_loss(pl)
        self._backward_reduce(pl)
        self._optimizer_step(pl)

    def _plan_for(self, n):
        """The launch plan for a batch of n <= N samples (built on first use)."""
        pl = self._plans.get(n)
        if pl is None:
            if not (0 < n < self.N):
                raise ValueError(f"batch of {n} samples: the trainer was built for {self.N} (smaller tails are allowed)")
            dev = self.device
            pl = dict(eng=Engine(self.eng.spec, n, dev, share_from=self.eng),
                      idx=torch.arange(n, dtype=torch.int64, device=dev),
                      ijf=torch.zeros(n, 3, dtype=torch.int8, device=dev),
                      tgt=torch.zeros(n, dtype=torch.int64, device=dev), graph=None, eager_first=True)
            self._plans[n] = pl
        return pl

    def _stage_inputs(self, idx, ijf, targets, images=None):
        n = len(ijf)
        pl = self._plan_for(n)
        if len(targets) != n or (idx is not None and len(idx) != n) or (images is not None and len(images) != n):
            raise ValueError("idx / ijf / targets / images must have the same length")
        s = self._slot
        self._slot = (s + 1) % self._ring
        if self._ring_events[s] is not None:
            self._ring_events[s].synchronize()
        self._pin_ijf[s].numpy()[:n] = ijf
        self._pin_tgt[s].numpy()[:n] = targets
        pl["ijf"].copy_(self._pin_ijf[s][:n], non_blocking=True)
        pl["tgt"].copy_(self._pin_tgt[s][:n], non_blocking=True)
        if images is not None:
            if self._pin_img is None:
                self._pin_img = [torch.zeros(self.N, 32, 32, 3, dtype=torch.uint8).pin_memory()
                                 for _ in range(self._ring)]
            self._pin_img[s].numpy()[:n] = images
            self.data[:n].copy_(self._pin_img[s][:n], non_blocking=True)
        else:
            self._pin_idx[s].numpy()[:n] = idx
            pl["idx"].copy_(self._pin_idx[s][:n], non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self._ring_events[s] = ev
        return pl

    @property
    def graph(self):
        return self._main["graph"]

    def _run(self, pl=None):
        pl = pl or self._main
        if not self.use_graph or pl.get("eager_first"):
            pl["eager_first"] = False        # a tail plan's first step runs eagerly (kernel attributes), then it is captured
            self._launch_step(pl)
        else:
            if pl["graph"] is None:
                self._capture(pl)
            pl["graph"].replay()
        if self.steps_done == 0:
            self.hyper_host[4] = 0.0
            self._push_hyper()
        self.steps_done += 1

    def _capture(self, pl):
        # warm-up outside capture is the caller's first eager step(s); here: capture only.
        # State mutated by a step (params, momentum, BN stats) is advanced by replays only, so save/restore
        # is unnecessary: capture itself does not execute kernels.
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._launch_step(pl)
        pl["graph"] = g

    def step(self, idx, ijf):
        """One training step on dataset samples `idx` (host int64 [n]) with crop/flip draws `ijf` (int8 [n,3]);
        n = N, or fewer for the ragged last batch of an epoch (its own plan and graph, built on first use)."""
        self._run(self._stage_inputs(idx, ijf, self.labels_host[idx]))
        return self.loss

    def step_host_batch(self, images_u8, labels, ijf):
        """One training step on a HOST uint8 batch [N,32,32,3] (copied to the device inside the call)."""
        assert not self.resident
        self._run(self._stage_inputs(None, ijf, labels, images=images_u8))
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
            ops.augment_im2col(ev["data"], ev["idx"], ev["ijf"], eng.x0, taps=eng.stem_taps)
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
        idx = hostdata.distributed_indices(n, world, rank, seed=seed, epoch=0)    # the reference never calls set_epoch
    else:
        idx = hostdata.randperm(hostdata.TorchCpuMT19937(seed), n)
    return hostdata.batches(idx, batch)
