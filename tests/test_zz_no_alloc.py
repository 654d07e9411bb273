"""GPU test of b2c_bn_bwd_stats_noalloc — the entry point for callers that cannot allocate device memory
around the call (an application whose GPU memory belongs to another allocator or a fixed budget, a
latency-bound loop that cannot take cudaMalloc's device synchronisation, a plug-in that only receives
input/output pointers). Its scratch is a fixed array inside libb2c.so instead of a caller buffer.

The check: through the C-ABI, its outputs equal those of the existing two-call form
(b2c_bn_bwd_reduce + b2c_bn_bwd_finalize with a caller-provided scratch) within the fp32 tolerance the
other kernel tests use, on two calls in a row with different shapes — the second call runs over whatever
the first left in the shared array. Oversized requests must be refused with B2C_ERR_TOO_LARGE.
This file sorts last on purpose, so a run that stops at the first failure has already run every other test.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

F32_TOL = 1e-3


def rel_err(a, b):
    a, b = a.double(), b.double()
    return ((a - b).norm() / b.norm().clamp_min(1e-30)).item()


def _case(ops, rows, C, seed, with_act):
    g = torch.Generator(device="cuda").manual_seed(seed)
    y = (torch.randn(rows, C, device="cuda", generator=g) * 1.5 + 0.3).to(torch.bfloat16)
    dout = torch.randn(rows, C, device="cuda", generator=g).to(torch.bfloat16)
    act = torch.randn(rows, C, device="cuda", generator=g).clamp_min(0).to(torch.bfloat16) if with_act else None
    gamma = torch.rand(C, device="cuda", generator=g) + 0.5
    mean = y.float().mean(0)
    rstd = (y.float().var(0, unbiased=False) + 1e-5).rsqrt()

    # existing entry points, caller-owned scratch
    prow = ops.bn_bwd_partial_rows(rows, C)
    partial = torch.empty(prow, 2, C, device="cuda")
    dz_ref = torch.empty_like(dout) if with_act else None
    ref = [torch.empty(C, device="cuda") for _ in range(5)]
    ops.bn_bwd_reduce(dout, act, y, dz_ref, rows, C, partial)
    ops.bn_bwd_finalize(partial, prow, C, rows, gamma, mean, rstd, *ref)

    # the allocation-free entry point: no scratch argument at all
    dz_new = torch.empty_like(dout) if with_act else None
    new = [torch.full((C,), float("nan"), device="cuda") for _ in range(5)]
    ops.bn_bwd_stats_noalloc(dout, act, y, dz_new, rows, C, rows, gamma, mean, rstd, *new)
    torch.cuda.synchronize()
    for name, a, b in zip(("dgamma", "dbeta", "coef_a", "coef_b", "coef_c"), new, ref):
        assert torch.isfinite(a).all(), name
        assert rel_err(a, b) < F32_TOL, (name, rows, C)
    if with_act:
        assert torch.equal(dz_new, dz_ref)


def test_noalloc_matches_two_call_form_twice_in_a_row():
    import b200cifar
    ops = b200cifar.ops
    _case(ops, rows=8192, C=512, seed=31, with_act=True)     # fills the shared array up to its limit
    _case(ops, rows=1000, C=64, seed=32, with_act=False)     # second call: smaller, over the stale contents
    _case(ops, rows=4096, C=128, seed=33, with_act=True)


def test_noalloc_rejects_sizes_beyond_its_scratch():
    import b200cifar
    from b200cifar import _lib
    ops = b200cifar.ops
    C = ops.NOALLOC_MAX_CHANNELS * 2
    rows = 64
    y = torch.zeros(rows, C, device="cuda", dtype=torch.bfloat16)
    v = [torch.zeros(C, device="cuda") for _ in range(8)]
    with pytest.raises(_lib.B2cError) as ei:
        ops.bn_bwd_stats_noalloc(y, None, y, None, rows, C, rows, *v)
    assert ei.value.code == _lib.B2C_ERR_TOO_LARGE
    assert "no-allocation" in str(ei.value)
