"""numpy restatement of the reference's optimiser and schedule (test oracle).

optim.SGD(lr, momentum=0.9, weight_decay=5e-4) as used at main.py:87-88 / main_dist.py:160-161:
dampening 0, no Nesterov, weight decay added to the gradient, first step initialises the buffer with the
(decayed) gradient. CosineAnnealingLR(T_max=200) at main.py:89 in closed form.
"""
import math

import numpy as np


def sgd_step(p, g, buf, lr, momentum=0.9, weight_decay=5e-4, first=False):
    """In-place on float32 numpy arrays; mirrors torch's single-tensor SGD arithmetic order."""
    d = (g + np.float32(weight_decay) * p).astype(np.float32)
    if first:
        buf[...] = d
    else:
        buf[...] = (np.float32(momentum) * buf + d).astype(np.float32)
    p[...] = (p - np.float32(lr) * buf).astype(np.float32)


def cosine_lr(base_lr, epoch, t_max=200, eta_min=0.0):
    return eta_min + (base_lr - eta_min) * (1.0 + math.cos(math.pi * epoch / t_max)) / 2.0
