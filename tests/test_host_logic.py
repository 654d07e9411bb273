"""CPU tests of the product's host-side logic (data ordering / draws, network spec, C-ABI surface)."""
import ctypes
import os
import re

import numpy as np
import torch

import b200cifar
from b200cifar import _lib, data, engine

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "tests", "golden")


def test_mt19937_matches_torch_generator():
    g = torch.Generator()
    g.manual_seed(2024)
    ref = torch.randint(0, 1000003, (2000,), generator=g, dtype=torch.int64).numpy()
    rng = data.TorchCpuMT19937(2024)
    got = rng.raw(2000).astype(np.int64) % 1000003
    assert np.array_equal(ref, got)


def test_draws_match_oracle_and_golden():
    from oracle import augment as oaug
    g = np.load(os.path.join(G, "augment_seed123.npz"))
    n = g["images"].shape[0]
    a = data.draw_crop_flip(data.TorchCpuMT19937(int(g["seed"])), n)
    b = oaug.draw_params(oaug.MT19937(int(g["seed"])), n)
    assert np.array_equal(a, b)
    big_a = data.draw_crop_flip(data.TorchCpuMT19937(9), 5000)
    big_b = oaug.draw_params(oaug.MT19937(9), 5000)
    assert np.array_equal(big_a, big_b)


def test_randperm_and_distributed_sampler():
    gen = torch.Generator()
    gen.manual_seed(0)
    assert np.array_equal(torch.randperm(50000, generator=gen).numpy(), data.randperm(data.TorchCpuMT19937(0), 50000))
    g = np.load(os.path.join(G, "sampler.npz"))
    for world in (1, 2, 4, 8):
        for rank in (0, world - 1):
            idx = data.distributed_indices(50000, world, rank)
            assert len(idx) == int(g[f"w{world}_r{rank}_len"])
            assert np.array_equal(idx[:64], g[f"w{world}_r{rank}_head"])
            assert int(idx.sum()) == int(g[f"w{world}_r{rank}_sum"])
    parts = data.batches(data.distributed_indices(50000, 8, 3), 128)
    assert len(parts) == 49 and len(parts[-1]) == 106         # 6250 = 48*128 + 106 (main_dist.py:111)


def test_resnet_spec_matches_reference_state_dict_names():
    g = np.load(os.path.join(G, "resnet18_seed0_n4.npz"))
    spec = engine.resnet_spec()
    assert list(spec["params"].keys()) == [str(s) for s in g["param_names"]]
    n = sum(int(np.prod(s)) for s in spec["params"].values())
    assert n == 11173962


def test_c_abi_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "b2c.h")).read()
    declared = set(re.findall(r"B2C_API\s+(?:const\s+)?\w+\*?\s+(b2c_\w+)\s*\(", hdr))
    assert len(declared) >= 20
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), name
    assert declared == set(_lib.SIGNATURES) | {"b2c_error_string"}
    l = _lib.load()
    assert l.b2c_version() >= 100
    assert b"not supported" in l.b2c_error_string(_lib.B2C_ERR_UNSUPPORTED)
    assert ctypes.sizeof(_lib.ConvDesc) == 52 and ctypes.sizeof(_lib.RepackEntry) == 48
