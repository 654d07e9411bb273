"""Explicit (no-autograd) restatement of the ResNet training step with selectable storage rounding.
TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Why it exists: the CUDA path stores activations and activation gradients in bf16 (fp32 accumulation and
statistics). Against the pure-fp32 oracle (oracle/nets.py) that storage costs ~0.2-0.4 % relative error
per layer in the forward pass and, because a ReLU mask flips wherever a pre-activation changes sign,
far larger differences in deep gradients — differences that come from the number format, not from the
kernels. This file restates the same step with the forward AND backward written out by hand
(models/resnet.py:47-51,132-140 and the textbook BatchNorm / ReLU / conv gradients), and a switch:

  round_bf16=False : pure fp32. Must agree with autograd on oracle/nets.py to ~1e-5 (tests/test_oracle.py);
                     this validates the hand-written backward.
  round_bf16=True  : rounds to bf16 exactly where the CUDA path stores bf16. The CUDA path must agree with
                     this to a tight tolerance (accumulation order only) — that isolates kernel bugs from
                     precision effects.
"""
from collections import OrderedDict

import torch
import torch.nn.functional as F

from .nets import RESNET18_BLOCKS, RESNET_PLANES

EPS = 1e-5
MOMENTUM = 0.1


def _r(t, on):
    return t.to(torch.bfloat16).to(torch.float32) if on else t


class _Unit:
    """conv + BatchNorm(train) with the tensors backward needs."""

    def __init__(self, name, bn, x, w, gamma, beta, stride, pad, rnd, buffers):
        self.name, self.bn, self.x, self.stride, self.pad, self.rnd = name, bn, x, stride, pad, rnd
        self.w_shape = w.shape
        self.wq = _r(w, rnd)
        y32 = F.conv2d(x, self.wq, stride=stride, padding=pad)
        n = y32.numel() / y32.shape[1]
        self.mean = y32.mean((0, 2, 3))
        self.var = y32.var((0, 2, 3), unbiased=False)          # statistics from the fp32 accumulators
        self.rstd = (self.var + EPS).rsqrt()
        self.y = _r(y32, rnd)                                   # stored conv output
        self.gamma = gamma
        self.scale = gamma * self.rstd
        self.shift = beta - self.mean * self.scale
        self.count = n
        if buffers is not None:
            rm, rv = buffers[bn + ".running_mean"], buffers[bn + ".running_var"]
            rm.mul_(1 - MOMENTUM).add_(MOMENTUM * self.mean)
            rv.mul_(1 - MOMENTUM).add_(MOMENTUM * self.var * n / (n - 1))

    def bn_out(self):
        return self.y * self.scale.view(1, -1, 1, 1) + self.shift.view(1, -1, 1, 1)

    def bn_backward(self, dz, grads):
        """dz: gradient w.r.t. BN output (already masked). Returns dy (stored rounding applied)."""
        s1 = dz.sum((0, 2, 3))
        s2 = (dz * self.y).sum((0, 2, 3))
        dbeta = s1
        dgamma = self.rstd * (s2 - self.mean * s1)
        grads[self.bn + ".weight"] = dgamma
        grads[self.bn + ".bias"] = dbeta
        n = self.count
        A = self.gamma * self.rstd
        B = -self.gamma * self.rstd * self.rstd * dgamma / n
        C = -self.gamma * self.rstd * dbeta / n + self.gamma * self.rstd * self.rstd * self.mean * dgamma / n
        dy = A.view(1, -1, 1, 1) * dz + (B.view(1, -1, 1, 1) * self.y + C.view(1, -1, 1, 1))
        return _r(dy, self.rnd)

    def wgrad(self, dy, grads):
        grads[self.name + ".weight"] = torch.nn.grad.conv2d_weight(self.x, self.w_shape, dy, stride=self.stride,
                                                                   padding=self.pad)

    def dgrad(self, dy):
        return torch.nn.grad.conv2d_input(self.x.shape, self.wq, dy, stride=self.stride, padding=self.pad)


def resnet_train_step(x, target, params, buffers, round_bf16, num_blocks=RESNET18_BLOCKS, taps=None, trace=None):
    """Returns (loss, logits, grads). `buffers` running stats are updated in place (pass None to skip).
    `trace` (optional dict) receives every stored tensor of the step, forward and backward, per block, so a
    test can teacher-force one block of the CUDA path at a time."""
    rnd = round_bf16
    P = params
    grads = OrderedDict()

    def tap(k, v):
        if taps is not None:
            taps[k] = v

    x0 = _r(x, rnd)
    stem = _Unit("conv1", "bn1", x0, P["conv1.weight"], P["bn1.weight"], P["bn1.bias"], 1, 1, rnd, buffers)
    tap("conv1", stem.y)
    a0 = _r(F.relu(stem.bn_out()), rnd)
    blocks = []
    out = a0
    in_planes = 64
    for li, (nb, planes) in enumerate(zip(num_blocks, RESNET_PLANES)):
        strides = [1 if li == 0 else 2] + [1] * (nb - 1)
        for bi, s in enumerate(strides):
            name = f"layer{li + 1}.{bi}"
            xin = out
            c1 = _Unit(name + ".conv1", name + ".bn1", xin, P[name + ".conv1.weight"], P[name + ".bn1.weight"],
                       P[name + ".bn1.bias"], s, 1, rnd, buffers)
            a1 = _r(F.relu(c1.bn_out()), rnd)
            c2 = _Unit(name + ".conv2", name + ".bn2", a1, P[name + ".conv2.weight"], P[name + ".bn2.weight"],
                       P[name + ".bn2.bias"], 1, 1, rnd, buffers)
            cs = None
            if (name + ".shortcut.0.weight") in P:
                cs = _Unit(name + ".shortcut.0", name + ".shortcut.1", xin, P[name + ".shortcut.0.weight"],
                           P[name + ".shortcut.1.weight"], P[name + ".shortcut.1.bias"], s, 0, rnd, buffers)
                pre = c2.bn_out() + cs.bn_out()
            else:
                pre = c2.bn_out() + xin
            out = _r(F.relu(pre), rnd)
            tap(name + ".conv1", c1.y); tap(name + ".conv2", c2.y); tap(name + ".out", out)
            blocks.append((c1, a1, c2, cs, out))
            in_planes = planes
    hw = out.shape[2] * out.shape[3]
    pooled = out.mean((2, 3))
    logits = F.linear(pooled, P["linear.weight"], P["linear.bias"])
    N = x.shape[0]
    logp = F.log_softmax(logits, 1)
    loss = -logp[torch.arange(N, device=x.device), target].mean()
    dlogits = (logp.exp() - F.one_hot(target, logits.shape[1]).float()) / N

    # ---------------- backward
    grads["linear.weight"] = dlogits.t() @ pooled
    grads["linear.bias"] = dlogits.sum(0)
    dpooled = dlogits @ P["linear.weight"]
    dout = _r((dpooled / hw).view(N, -1, 1, 1).expand_as(out).contiguous(), rnd)
    if trace is not None:
        trace.update(stem=stem, a0=a0, blocks=blocks, pooled=pooled, dlogits=dlogits, dout_in=[], dx_out=[])
    for (c1, a1, c2, cs, out_b) in reversed(blocks):
        if trace is not None:
            trace["dout_in"].append(dout)
        dz = dout * (out_b > 0).float()
        dy2 = c2.bn_backward(dz, grads)
        dys = cs.bn_backward(dz, grads) if cs is not None else None
        c2.wgrad(dy2, grads)
        da1 = _r(c2.dgrad(dy2), rnd)
        dz1 = da1 * (a1 > 0).float()
        dy1 = c1.bn_backward(dz1, grads)
        c1.wgrad(dy1, grads)
        if cs is not None:
            cs.wgrad(dys, grads)
            dx0 = _r(cs.dgrad(dys), rnd)
            dout = _r(c1.dgrad(dy1) + dx0, rnd)
        else:
            dout = _r(c1.dgrad(dy1) + dz, rnd)
        if trace is not None:
            trace["dx_out"].append(dout)
    dz0 = dout * (a0 > 0).float()
    dy0 = stem.bn_backward(dz0, grads)
    stem.wgrad(dy0, grads)
    ordered = OrderedDict((k, grads[k]) for k in params.keys())
    return loss, logits, ordered
