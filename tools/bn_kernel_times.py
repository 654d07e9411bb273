"""Time the BN elementwise kernels alone at the ResNet-18 batch-1024 activation sizes (HBM-bound: report GB/s of
algorithmic bytes). Run on the GPU box: python tools/bn_kernel_times.py > gpurun_out/bn_kernel_times.json"""
import json
import sys

import torch

sys.path.insert(0, ".")
from b200cifar import ops  # noqa: E402


def timeit(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3   # us


def main():
    out = []
    for rows, C in [(1024 * 32 * 32, 64), (1024 * 16 * 16, 128), (1024 * 8 * 8, 256), (1024 * 4 * 4, 512)]:
        g = torch.Generator(device="cuda").manual_seed(0)
        mk = lambda: torch.randn(rows, C, device="cuda", generator=g).to(torch.bfloat16)
        y, dout, act, res, dz, dy = mk(), mk(), mk(), mk(), mk(), mk()
        vec = lambda: torch.rand(C, device="cuda", generator=g) + 0.5
        scale, shift, ca, cb, cc = vec(), vec() - 1.0, vec(), vec(), vec()
        prow = ops.bn_bwd_partial_rows(rows, C)
        partial = torch.empty(prow, 2, C, device="cuda")
        elt = rows * C
        cases = {
            "act_apply": (lambda: ops.bn_act_apply(y, scale, shift, act, rows, C, relu=True), 4),
            "act_apply_res": (lambda: ops.bn_act_apply(y, scale, shift, act, rows, C, relu=True, res=res), 6),
            "bwd_reduce_act_dz": (lambda: ops.bn_bwd_reduce(dout, act, y, dz, rows, C, partial), 8),
            "bwd_reduce_masky": (lambda: ops.bn_bwd_reduce(dout, None, y, None, rows, C, partial, scale, shift), 4),
            "bwd_reduce_plain": (lambda: ops.bn_bwd_reduce(dout, None, y, None, rows, C, partial), 4),
            "bwd_apply": (lambda: ops.bn_bwd_apply(dz, y, ca, cb, cc, dy, rows, C), 6),
            "bwd_apply_masky": (lambda: ops.bn_bwd_apply(dz, y, ca, cb, cc, dy, rows, C, relu_scale=scale,
                                                         relu_shift=shift), 6),
            "bwd_apply_add": (lambda: ops.bn_bwd_apply(dz, y, ca, cb, cc, dy, rows, C, add=res), 8),
        }
        for name, (fn, bpe) in cases.items():
            us = timeit(fn)
            out.append({"rows": rows, "C": C, "kernel": name, "us": round(us, 1),
                        "GBps": round(elt * bpe / us / 1e3, 0)})
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
