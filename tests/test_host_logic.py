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
    assert ctypes.sizeof(_lib.ConvDesc) == 60 and ctypes.sizeof(_lib.RepackEntry) == 48


def test_spec_parameter_counts_match_the_reference():
    """parameter totals of every supported family (the arenas are laid out from them)."""
    want = {"resnet_spec": 11173962, "vgg_spec": 14728266, "preact_resnet_spec": 11171146,
            "mobilenetv2_spec": 2296922, "resnet50_spec": 23520842}
    for fn, n in want.items():
        spec = getattr(engine, fn)()
        assert sum(int(np.prod(s)) for s in spec["params"].values()) == n, fn
        off, total = engine.param_offsets(spec)
        # offsets are 16-byte aligned, in parameter order, non-overlapping
        prev_end = 0
        for name, (o, cnt, shape) in off.items():
            assert o % 4 == 0 and o >= prev_end and cnt == int(np.prod(shape)), name
            prev_end = o + cnt
        assert total >= prev_end


def test_c_abi_rejects_bad_arguments_without_touching_the_gpu():
    """Error convention of include/b2c.h: every entry point validates before its first CUDA call, so these return
    codes can be checked on a machine without a GPU. Nothing here launches a kernel."""
    l = _lib.load()
    fake = ctypes.c_void_p(0x10000)                     # aligned, never dereferenced on these paths
    odd = ctypes.c_void_p(0x10008)                      # not 16-byte aligned
    rows = ctypes.c_int(0)

    def desc(**kw):
        base = dict(n=4, h=32, w=32, cin=64, cout=64, ksize=3, stride=1, cin_valid=0, cout_valid=0, kflat_cin=0,
                    block_n=0, max_ctas=0, flags=0, x_ld=0, y_ld=0)
        base.update(kw)
        return _lib.ConvDesc(*[base[f] for f, _ in _lib.ConvDesc._fields_])

    raw = l.raw if hasattr(l, "raw") else (lambda name: getattr(l, name))
    fprop, dgrad, wgrad = raw("b2c_conv_fprop"), raw("b2c_conv_dgrad"), raw("b2c_conv_wgrad")
    U, A = _lib.B2C_ERR_UNSUPPORTED, _lib.B2C_ERR_ARG
    assert fprop(None, fake, fake, fake, None, None, None, ctypes.byref(rows), None) == A
    for bad in (dict(x_ld=32), dict(y_ld=68)):          # a pitch below the channel count / off the 16-byte grid
        assert fprop(ctypes.byref(desc(**bad)), fake, fake, fake, None, None, None, ctypes.byref(rows), None) == A, bad
    for bad in (dict(cin=48), dict(cout=96 + 8), dict(ksize=5), dict(stride=3), dict(stride=2, h=31)):
        assert fprop(ctypes.byref(desc(**bad)), fake, fake, fake, None, None, None, ctypes.byref(rows), None) == U, bad
    assert fprop(ctypes.byref(desc(block_n=96)), fake, fake, fake, None, None, None, ctypes.byref(rows), None) == A
    assert fprop(ctypes.byref(desc()), None, fake, fake, None, None, None, ctypes.byref(rows), None) == A
    assert fprop(ctypes.byref(desc()), odd, fake, fake, None, None, None, ctypes.byref(rows), None) == A
    assert fprop(ctypes.byref(desc()), fake, fake, fake, None, fake, None, ctypes.byref(rows), None) == A   # sum w/o sq
    assert dgrad(ctypes.byref(desc()), fake, fake, None, None, None) == A
    assert wgrad(ctypes.byref(desc()), fake, None, fake, None, 0, None) == A
    assert raw("b2c_conv_dgrad_with_shortcut")(ctypes.byref(desc()), fake, fake, fake, fake, None, None) == U  # s1
    assert raw("b2c_conv_dgrad_bn")(ctypes.byref(desc()), fake, fake, fake, None, None, fake, fake, fake,
                                    ctypes.byref(rows), None) == A
    assert raw("b2c_sgd_step")(fake, fake, fake, 6, fake, None) == A                      # n % 4 != 0
    assert raw("b2c_sgd_step")(fake, fake, fake, 0, fake, None) == 0                      # empty arena: no launch
    assert raw("b2c_bn_finalize")(fake, fake, 0, 64, 1.0, fake, fake, None, 1e-5, 0.1, None, None, fake, fake, fake,
                                  fake, None) == A                                        # rows < 1
    assert raw("b2c_bn_bwd_partial_rows")(1 << 20, 64) == 148 * 4
    assert raw("b2c_bn_bwd_partial_rows")(10, 64) == 1
    assert raw("b2c_bn_bwd_partial_rows")(10, 12) == U
    # entry points added for the wider model zoo: shapes off the vector grid, pitches below the channel count, ...
    assert raw("b2c_maxpool3x3_fwd")(fake, fake, fake, 4, 32, 32, 64, 3, None) == U               # stride 3
    assert raw("b2c_maxpool3x3_fwd")(fake, fake, fake, 4, 32, 32, 60, 1, None) == U               # channels % 8
    assert raw("b2c_maxpool3x3_bwd")(fake, None, None, fake, 4, 32, 32, 64, 1, None) == A         # no argmax
    assert raw("b2c_copy_channels")(fake, 64, 32, fake, 64, 0, 10, 64, None) == U                 # slice past the pitch
    assert raw("b2c_copy_channels")(fake, 64, 0, odd, 64, 0, 10, 64, None) == A
    assert raw("b2c_grouped_wgrad_extract")(fake, fake, 128, 128, 9, 3, None) == A                # 128 % 3
    assert raw("b2c_head_fwd_ld")(fake, 4, 64, 368, 360, fake, fake, 10, fake, fake, None) == U   # pitch < channels
    assert raw("b2c_se_fwd")(fake, 4, 64, 128, 100, 8, fake, fake, fake, fake, fake, fake, fake, None, fake, None) == U
    assert raw("b2c_se_bwd")(fake, None, 4, 64, 128, 128, 8, *([fake] * 12), fake, None) == A     # no y
    d1 = desc(ksize=1, cin=128)
    ptrs = (ctypes.c_void_p * 4)(0x10000, 0x10000, 0x10000, 0x10000)
    cat_n, cat_r = raw("b2c_conv1x1_catn_fprop"), raw("b2c_conv1x1_catr_fprop")
    assert cat_n(ctypes.byref(d1), ptrs, 5, fake, fake, None, None, ctypes.byref(rows), None) == U      # > 4 sources
    assert cat_n(ctypes.byref(desc(cin=128)), ptrs, 4, fake, fake, None, None, ctypes.byref(rows), None) == U   # 3x3
    cins = (ctypes.c_int * 4)(64, 192, 64, 64)
    assert cat_r(ctypes.byref(d1), ptrs, cins, 4, fake, fake, None, None, ctypes.byref(rows), None) == U   # wider than d.cin
    cins = (ctypes.c_int * 4)(64, 96, 64, 64)
    assert cat_r(ctypes.byref(d1), ptrs, cins, 4, fake, fake, None, None, ctypes.byref(rows), None) == U   # 96 % 64
    assert cat_r(ctypes.byref(d1), ptrs, None, 4, fake, fake, None, None, ctypes.byref(rows), None) == A
    noalloc = raw("b2c_bn_bwd_stats_noalloc")
    args = [fake, None, fake, None, 1024, 1024, 1024.0] + [fake] * 8 + [None]
    assert noalloc(*args) == _lib.B2C_ERR_TOO_LARGE
    for code, word in ((U, b"not supported"), (A, b"invalid argument"), (_lib.B2C_ERR_TOO_LARGE, b"exceeds"),
                       (0, b"success")):
        assert word in l.b2c_error_string(code)


def test_package_initialiser_matches_reference_order():
    """engine.init_params reproduces `torch.manual_seed(s); <reference constructor>()` bit for bit: checked against the
    oracle's builders, which tests/test_oracle.py pins to fixtures generated by the reference's own modules."""
    import torch
    from b200cifar import engine
    from oracle import nets
    cases = [(engine.resnet_spec(), nets.build_resnet_params), (engine.vgg_spec("VGG16"), nets.build_vgg_params),
             (engine.preact_resnet_spec(), nets.build_preact_params),
             (engine.mobilenetv2_spec(), nets.build_mobilenetv2_params),
             (engine.resnet50_spec(), nets.build_bottleneck_params),
             (engine.resnext_spec(), nets.build_resnext_params),
             (engine.resnext_spec(cardinality=32, bottleneck_width=4), nets.build_resnext_32x4d_params),
             (engine.resnext29_4x64d_spec(), nets.build_resnext_4x64d_params),
             (engine.regnetx_spec("RegNetX_200MF"), nets.build_regnetx_params),
             (engine.regnetx_spec("RegNetX_400MF"), nets.build_regnetx_400mf_params),
             (engine.dla_spec(), nets.build_dla_params), (engine.simple_dla_spec(), nets.build_simpledla_params),
             (engine.senet_spec(), nets.build_senet_params), (engine.regnety_400mf_spec(), nets.build_regnety_400mf_params),
             (engine.googlenet_spec(), nets.build_googlenet_params),
             (engine.preact_resnet50_spec(), nets.build_preact_bottleneck_params),
             (engine.mobilenet_spec(), nets.build_mobilenet_params)]
    for spec, build in cases:
        for seed in (0, 3):
            p, b = engine.init_params(spec, seed)
            rp, rb = build(seed)
            assert list(p) == list(rp), spec["family"]
            assert all(torch.equal(p[k], rp[k]) for k in p), spec["family"]
            assert all(torch.equal(b[k], rb[k]) for k in b), spec["family"]


def test_epoch_plan_batches_and_seed():
    """make_epoch_plan: DataLoader batching with the ragged tail kept (main.py:45, main_dist.py:111-127)."""
    from b200cifar import data, trainer
    plan = trainer.make_epoch_plan(50000, 128)
    assert len(plan) == 391 and len(plan[-1]) == 80
    d0 = trainer.make_epoch_plan(50000, 128, world=8, rank=3, seed=0, distributed=True)
    assert len(d0) == 49 and len(d0[-1]) == 106
    assert np.array_equal(np.concatenate(d0), data.distributed_indices(50000, 8, 3, seed=0))
    d1 = trainer.make_epoch_plan(50000, 128, world=8, rank=3, seed=1, distributed=True)
    assert not np.array_equal(np.concatenate(d0), np.concatenate(d1))


def test_bench_algorithmic_bytes():
    """bench.py's HBM-roofline numerator for MobileNetV2 = 3 x the fused-minimum forward traffic of SURVEY.md §8a
    (880 MB per 128 images)."""
    import bench
    from b200cifar import engine
    per_img = bench.algorithmic_bytes_per_image(engine.mobilenetv2_spec(), engine)
    assert abs(per_img * 128 / 3 / 1e6 - 880) < 5
    assert abs(bench.algorithmic_bytes_per_image(engine.resnet_spec(), engine) * 128 / 3 / 1e6 - 327.9) < 1


def test_spec_from_the_references_own_modules():
    """accelerate.spec_from_module on modules built by the reference's OWN model files (the tree itself, or the copy
    staged under oracle/_ref): the derived spec lists exactly the module's parameters, names, order and shapes, for
    every supported constructor; an unsupported one raises instead of falling back."""
    import importlib
    import sys
    import pytest
    from b200cifar.accelerate import spec_from_module
    from oracle.stage_reference import reference_dir
    ref = reference_dir()
    if ref is None:
        pytest.skip("reference files neither present nor staged")
    if ref not in sys.path:
        sys.path.insert(0, ref)
    cases = [("resnet", n, ()) for n in ("ResNet18", "ResNet34", "ResNet50", "ResNet101", "ResNet152")] + \
            [("vgg", "VGG", (v,)) for v in ("VGG11", "VGG13", "VGG16", "VGG19")] + \
            [("preact_resnet", "PreActResNet18", ()), ("preact_resnet", "PreActResNet34", ()),
             ("preact_resnet", "PreActResNet50", ()), ("preact_resnet", "PreActResNet101", ()),
             ("preact_resnet", "PreActResNet152", ()),
             ("mobilenetv2", "MobileNetV2", ()), ("mobilenet", "MobileNet", ()), ("dla_simple", "SimpleDLA", ())] + \
            [("resnext", n, ()) for n in ("ResNeXt29_2x64d", "ResNeXt29_4x64d", "ResNeXt29_8x64d", "ResNeXt29_32x4d")] + \
            [("regnet", "RegNetX_200MF", ()), ("regnet", "RegNetX_400MF", ()), ("dla", "DLA", ()),
             ("senet", "SENet18", ()), ("regnet", "RegNetY_400MF", ()),
             ("googlenet", "GoogLeNet", ())]
    for mod, ctor, args in cases:
        net = getattr(importlib.import_module("models." + mod), ctor)(*args)
        spec = spec_from_module(net)
        want = [(n, tuple(p.shape)) for n, p in net.named_parameters()]
        got = [(n, tuple(s)) for n, s in spec["params"].items()]
        assert got == want, (mod, ctor)
        bufs = {n for n, _ in net.named_buffers() if not n.endswith("num_batches_tracked")}
        assert set(spec["buffers"]) == bufs, (mod, ctor)
    for mod, ctor in (("dpn", "DPN92"), ("densenet", "DenseNet121"), ("lenet", "LeNet"), ("efficientnet", "EfficientNetB0")):
        with pytest.raises(NotImplementedError):
            spec_from_module(getattr(importlib.import_module("models." + mod), ctor)())
