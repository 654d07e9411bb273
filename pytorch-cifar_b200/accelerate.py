"""Drop-in surface: keep the reference's nn.Module / autograd / optimizer API, run the B200 kernels underneath.

    net = ResNet18().cuda()                 # the reference's unmodified models/resnet.py
    net = accelerate(net)                   # same object: same parameters, state_dict keys, train()/eval()
    outputs = net(inputs)                   # main.py:102  — fp32 NCHW in, fp32 [N,10] logits out
    loss = criterion(outputs, targets); loss.backward(); optimizer.step()      # main.py:103-105

What changes: `net.forward` becomes one torch.autograd.Function whose forward/backward run the static kernel
plan of engine.py; every nn.Parameter / BN buffer becomes a view into the engine's flat fp32 arenas (same
names, shapes, dtypes, so checkpoints round-trip: main.py:77-84,138-148); `.grad` tensors handed to autograd
are views of the gradient arena, so optim.SGD, DDP's reducer hooks and GradScaler-free scripts work unchanged.
Supported compositions: ResNet with BasicBlock or Bottleneck (models/resnet.py), VGG (models/vgg.py), PreActResNet with
PreActBlock (models/preact_resnet.py), MobileNetV2 (models/mobilenetv2.py), MobileNet (models/mobilenet.py), SimpleDLA (models/dla_simple.py), ResNeXt
(models/resnext.py, grouped convs), RegNetX / RegNetY (models/regnet.py), DLA (models/dla.py), SENet18 (models/senet.py), GoogLeNet (models/googlenet.py); anything else raises — there is no fallback path.

Recognition is structural (attribute names and layer shapes), not by import path: `from models import *`
shadows class names across files and the reference tree may sit anywhere on sys.path.
"""
import threading

import torch
import torch.nn as nn

from . import ops
from .engine import (Engine, dla_spec, googlenet_spec, mobilenet_spec, mobilenetv2_spec, plan_grad_buckets, preact_resnet_spec, regnetx_spec, resnet_spec,
                     resnext_spec, senet_spec, simple_dla_spec, vgg_spec, REGNET_CFG, VGG_CFG)


def _is(m, name):
    return type(m).__name__ == name


def spec_from_module(net):
    """Derive the engine spec from a reference-style module by structure."""
    if _is(net, "ResNet") and hasattr(net, "layer4") and hasattr(net, "linear") and hasattr(net, "bn1"):
        blocks = [len(getattr(net, f"layer{i}")) for i in range(1, 5)]
        kinds = {type(b).__name__ for i in range(1, 5) for b in getattr(net, f"layer{i}")}
        if kinds == {"BasicBlock"}:
            return resnet_spec(tuple(blocks), num_classes=net.linear.out_features)
        if kinds == {"Bottleneck"}:                       # ResNet-50/101/152: what main_dist.py:136 hard-codes
            return resnet_spec(tuple(blocks), num_classes=net.linear.out_features, bottleneck=True)
        raise NotImplementedError(f"accelerate: unsupported ResNet block types {sorted(kinds)}")
    if _is(net, "ResNeXt") and hasattr(net, "layer3") and not hasattr(net, "layer4") and hasattr(net, "cardinality"):
        blocks = [len(getattr(net, f"layer{i}")) for i in range(1, 4)]       # models/resnext.py:48-52: three stages
        conv2 = net.layer1[0].conv2
        return resnext_spec(tuple(blocks), cardinality=conv2.groups, bottleneck_width=conv2.out_channels // conv2.groups,
                            num_classes=net.linear.out_features)
    if _is(net, "RegNet") and hasattr(net, "cfg") and hasattr(net, "layer4"):
        cfg = net.cfg                                                         # models/regnet.py:68-71
        for name, ref in REGNET_CFG.items():
            if cfg.get("se_ratio", 0) == ref["se_ratio"] and cfg.get("bottleneck_ratio") == 1 and \
                    all(tuple(cfg[k]) == ref[k] for k in ("depths", "widths", "strides")) and \
                    cfg["group_width"] == ref["group_width"]:
                return regnetx_spec(name, num_classes=net.linear.out_features)
        raise NotImplementedError("accelerate: only the RegNet configurations of models/regnet.py are supported")
    if _is(net, "SENet") and hasattr(net, "layer4") and hasattr(net, "bn1"):
        blocks = [len(getattr(net, f"layer{i}")) for i in range(1, 5)]
        if not all(_is(b, "PreActBlock") and hasattr(b, "fc1") for i in range(1, 5) for b in getattr(net, f"layer{i}")):
            raise NotImplementedError("accelerate: only SENets of PreActBlocks (SENet18) are supported")
        return senet_spec(tuple(blocks), num_classes=net.linear.out_features)
    if _is(net, "GoogLeNet") and hasattr(net, "pre_layers") and hasattr(net, "b5") and _is(net.a3, "Inception"):
        return googlenet_spec(num_classes=net.linear.out_features)
    if _is(net, "PreActResNet") and hasattr(net, "layer4"):
        blocks = [len(getattr(net, f"layer{i}")) for i in range(1, 5)]
        kinds = {type(b).__name__ for i in range(1, 5) for b in getattr(net, f"layer{i}")}
        if kinds == {"PreActBlock"}:
            return preact_resnet_spec(tuple(blocks), num_classes=net.linear.out_features)
        if kinds == {"PreActBottleneck"}:                 # PreActResNet-50/101/152
            return preact_resnet_spec(tuple(blocks), num_classes=net.linear.out_features, bottleneck=True)
        raise NotImplementedError(f"accelerate: unsupported PreActResNet block types {sorted(kinds)}")
    if _is(net, "VGG") and hasattr(net, "features") and hasattr(net, "classifier"):
        cfg = []
        for m in net.features:
            if isinstance(m, nn.Conv2d):
                cfg.append(m.out_channels)
            elif isinstance(m, nn.MaxPool2d):
                cfg.append("M")
        for name, ref in VGG_CFG.items():
            if ref == cfg:
                return vgg_spec(name, num_classes=net.classifier.out_features)
    if _is(net, "SimpleDLA") and hasattr(net, "layer6") and hasattr(net, "base") and _is(net.layer3, "Tree"):
        return simple_dla_spec(num_classes=net.linear.out_features)       # what main.py:71 constructs
    if _is(net, "DLA") and hasattr(net, "layer6") and hasattr(net, "base") and hasattr(net.layer4, "level_1") \
            and hasattr(net.layer4, "prev_root"):
        return dla_spec(num_classes=net.linear.out_features)              # models/dla.py
    if _is(net, "MobileNet") and hasattr(net, "layers") and len(net.layers) == 13 and hasattr(net, "bn1"):
        return mobilenet_spec(num_classes=net.linear.out_features)
    if _is(net, "MobileNetV2") and hasattr(net, "layers") and hasattr(net, "conv2") and len(net.layers) == 17:
        return mobilenetv2_spec(num_classes=net.linear.out_features)
    raise NotImplementedError(f"accelerate: unsupported module {type(net).__module__}.{type(net).__name__}")


class _PlanFn(torch.autograd.Function):
    """forward/backward of the whole network as one autograd node (parameters are explicit inputs so that
    autograd — and DDP's hooks on AccumulateGrad — see a gradient for each of them)."""

    @staticmethod
    def forward(ctx, state, x, *params):
        eng = state.engine_for(x.shape[0])
        ctx.state, ctx.eng = state, eng
        training = state.net.training

        def launch(xin):
            eng.repack()                                 # parameters may have changed (optimizer, load_state_dict)
            ops.stem_im2col(xin, eng.x0, taps=eng.stem_taps)
            eng.forward(training=training)
        if state.use_graphs:
            xin = state.static_input(eng)
            xin.copy_(x)
            state.replay(("fwd", eng.N, training), lambda: launch(xin))
        else:
            launch(x.contiguous().float())
        if training:
            state.bump_batches_tracked()
        return eng.logits.clone()

    @staticmethod
    def backward(ctx, dlogits):
        eng, state = ctx.eng, ctx.state
        eng.dlogits.copy_(dlogits)
        # The tensors returned below are views of the gradient arena G. When a parameter has no .grad yet,
        # AccumulateGrad adopts the view as .grad (no copy) — the fast path, and what zero_grad(set_to_none=True)
        # (main.py:101) gives every step. If a .grad that ALIASES G survives into this backward
        # (zero_grad(set_to_none=False), p.grad.zero_(), gradient accumulation, two backwards before a step),
        # recomputing into G would wipe it and AccumulateGrad's `.grad += view` would then add the new gradient
        # to itself. In that case the new gradient is computed in G, moved to a private copy, G is restored, and
        # autograd accumulates the copy into the surviving .grad (= G) as usual.
        g_store = eng.G.untyped_storage().data_ptr()
        aliased = any(p.grad is not None and p.grad.untyped_storage().data_ptr() == g_store for p in state.params)
        keep = eng.G.clone() if aliased else None

        def launch():
            eng.zero_grad()
            if state.ddp is not None:
                d = state.ddp
                eng.backward_allreduce(d["buckets"], d["ready_after"], d["comm_stream"], group=d["group"], average=True)
            else:
                eng.backward()
        if state.use_graphs and state.ddp is None:
            state.replay(("bwd", eng.N), launch)
        else:
            launch()
        if aliased:
            fresh = eng.G.clone()
            eng.G.copy_(keep)
            grads = [eng._arena_view(fresh, name) for name in state.param_names]
        else:
            grads = [eng.grad_view(name) for name in state.param_names]
        return (None, None, *grads)


class _State:
    def __init__(self, net, spec):
        self.net, self.spec = net, spec
        self.engines = {}
        self.param_names = list(spec["params"].keys())
        self.named = dict(net.named_parameters())
        missing = [n for n in self.param_names if n not in self.named]
        if missing or len(self.named) != len(self.param_names):
            raise NotImplementedError(f"accelerate: parameter names do not match the reference layout: {missing[:3]}")
        self.params = [self.named[n] for n in self.param_names]
        self.nbt = [b for n, b in net.named_buffers() if n.endswith("num_batches_tracked")]
        self.ddp = None                      # set by attach_ddp(): gradient buckets are all-reduced inside backward
        self.lock = threading.Lock()         # one forward at a time: the plan's buffers are not re-entrant
        self.device = self.params[0].device
        # The arenas exist from here on (a minimal plan owns them; per-batch-size plans share them), and the
        # module's parameters / BN buffers are views of them BEFORE anything else looks at the module — in
        # particular DistributedDataParallel's constructor (main_dist.py:141-144), whose Reducer records parameter
        # strides and broadcasts parameter storage.
        self.base = Engine(spec, 2, self.device)
        self.engines[2] = self.base
        self._adopt(self.base)
        # accelerate(net, graphs=True): the launch sequences of forward and of backward (static buffers, static order)
        # are captured into CUDA graphs per (batch size, mode) on their second use and replayed afterwards
        self.use_graphs = False
        self._graphs, self._seen, self._xin = {}, set(), {}

    def static_input(self, eng):
        t = self._xin.get(eng.N)
        if t is None:
            t = self._xin[eng.N] = torch.zeros(eng.N, 3, 32, 32, device=self.device, dtype=torch.float32)
        return t

    def replay(self, key, launch):
        """First use of `key`: run eagerly (kernel attributes, NCCL, allocator warm-up). Second use: capture, then
        replay. Later uses: replay."""
        g = self._graphs.get(key)
        if g is None:
            if key not in self._seen:
                self._seen.add(key)
                launch()
                return
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, capture_error_mode="thread_local"):
                launch()
            self._graphs[key] = g
        g.replay()

    def engine_for(self, batch):
        eng = self.engines.get(batch)
        if eng is None:
            eng = Engine(self.spec, batch, self.device, share_from=self.base)
            self.engines[batch] = eng
        return eng

    def _adopt(self, eng):
        """Move parameters and BN running statistics into the arenas and re-point the module at them."""
        with torch.no_grad():
            for name, p in zip(self.param_names, self.params):
                view = eng.param_view(name)
                view.copy_(p.detach().to(view.dtype))
                p.data = view
            for name, b in self.net.named_buffers():
                if name in eng.b_off:
                    view = eng.buffer_view(name)
                    view.copy_(b.detach().to(view.dtype))
                    mod, attr = name.rsplit(".", 1)
                    self.net.get_submodule(mod)._buffers[attr] = view

    def bump_batches_tracked(self):
        if self.nbt:
            torch._foreach_add_(self.nbt, 1)              # nn.BatchNorm2d.num_batches_tracked += 1


def accelerate(net, graphs=False):
    """Replace net.forward with the B200 plan. Call after `net.to('cuda')`; returns the same object.
    graphs=True replays the forward and the backward launch sequences as CUDA graphs (one per batch size and mode,
    captured on second use): the ~160 launches of a step are issued eagerly from Python otherwise, which bounds small
    batches by the host (batch 128: 37 k images/s eager against 107 k for the graph-replayed Trainer). With
    attach_ddp() the backward, which contains the NCCL calls, stays eager."""
    if not next(net.parameters()).is_cuda:
        raise RuntimeError("accelerate(net): move the network to a CUDA device first (there is no CPU path)")
    spec = spec_from_module(net)
    state = _State(net, spec)
    state.use_graphs = bool(graphs)

    def forward(x):
        if x.dim() != 4 or tuple(x.shape[1:]) != (3, 32, 32):
            raise ValueError(f"accelerate: expected [N,3,32,32] input, got {tuple(x.shape)}")
        if x.device != state.device:
            raise RuntimeError(
                f"accelerate: input on {x.device} but the network's arenas live on {state.device}. An accelerated "
                "network is bound to ONE GPU: nn.DataParallel over several GPUs (main.py:73-75, main_dist.py without "
                "--dist) replicates the module and would drive this plan from every replica thread. Run with one "
                "visible GPU (CUDA_VISIBLE_DEVICES=0) or use DistributedDataParallel, one process per GPU "
                "(main_dist.py --dist).")
        if not state.lock.acquire(blocking=False):
            raise RuntimeError("accelerate: forward() re-entered from another thread (nn.DataParallel replicas?); "
                               "the kernel plan owns one set of buffers and runs one forward at a time")
        try:
            if torch.is_grad_enabled() and net.training:
                return _PlanFn.apply(state, x, *state.params)
            with torch.no_grad():
                return _PlanFn.forward(_Ctx(), state, x)
        finally:
            state.lock.release()
    net.forward = forward
    net._b2c_state = state
    return net


class _Ctx:
    """Stand-in ctx for the no-grad path."""
    pass


def _ddp_passthrough_hook(_state, bucket):
    """DDP communication hook of an attached network: the bucket already holds averaged gradients (they were
    all-reduced as arena buckets while backward was still running), so the Reducer has nothing left to send."""
    fut = torch.futures.Future()
    fut.set_result(bucket.buffer())
    return fut


def attach_ddp(ddp_net, bucket_bytes=25 * 2 ** 20):
    """Call on `DistributedDataParallel(accelerate(net), device_ids=[gpu])` (main_dist.py:141-144). Without it
    DDP works as for any module, but the whole network is ONE autograd node, so the Reducer sees every gradient
    at the same instant — at the end of backward — and its all-reduce overlaps nothing. After attach_ddp the
    gradient arena is all-reduced (NCCL average over ddp_net.process_group) in buckets on a side stream from
    INSIDE the backward plan, each bucket as soon as its last weight gradient is enqueued (what the Reducer does
    for a layer-by-layer autograd graph), and a communication hook tells the Reducer the buckets are done.
    Returns ddp_net."""
    import torch.distributed as dist
    net = ddp_net.module
    state = getattr(net, "_b2c_state", None)
    if state is None:
        raise ValueError("attach_ddp needs DistributedDataParallel around a network returned by accelerate()")
    group = ddp_net.process_group
    if dist.get_backend(group) != "nccl":
        raise RuntimeError("attach_ddp: the process group must use the nccl backend")
    buckets, ready_after = plan_grad_buckets(state.spec, bucket_bytes)
    state.ddp = dict(group=group, buckets=buckets, ready_after=ready_after,
                     comm_stream=torch.cuda.Stream(device=state.device))
    ddp_net.register_comm_hook(None, _ddp_passthrough_hook)
    return ddp_net


class FusedSGD(torch.optim.Optimizer):
    """optim.SGD(params, lr, momentum, weight_decay) (main.py:87-88) as ONE kernel over the parameter arena of an
    accelerated network. Same constructor arguments and param_groups (so lr schedulers work); dampening 0, no
    Nesterov, as the reference uses it."""

    def __init__(self, net, lr=0.1, momentum=0.9, weight_decay=5e-4):
        state = getattr(net, "_b2c_state", None)
        if state is None:
            raise ValueError("FusedSGD needs a network returned by accelerate()")
        super().__init__(list(net.parameters()), dict(lr=lr, momentum=momentum, weight_decay=weight_decay))
        self._state = state
        self._hyper = None
        self._steps = 0

    @torch.no_grad()
    def step(self, closure=None):
        eng = self._state.base
        if eng is None:
            raise RuntimeError("FusedSGD.step() before the first forward/backward")
        g = self.param_groups[0]
        vals = [g["lr"], g["momentum"], g["weight_decay"], 1.0, 1.0 if self._steps == 0 else 0.0]
        if self._hyper is None:
            self._hyper = torch.tensor(vals, dtype=torch.float32, device=eng.device)
        else:
            self._hyper.copy_(torch.tensor(vals, dtype=torch.float32), non_blocking=True)
        ops.sgd_step(eng.P, eng.G, eng.M, self._hyper)
        self._steps += 1

    def zero_grad(self, set_to_none=True):
        super().zero_grad(set_to_none=set_to_none)
