"""world_size-2 CPU (gloo) test of the data-parallel host logic: sample sharding, gradient bucketing and the
bucketed all-reduce + 1/world averaging that trainer.py performs with NCCL on the GPU
(reference: DistributedSampler main_dist.py:110, DistributedDataParallel main_dist.py:141-144)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import b200cifar
from b200cifar import data, engine


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    spec = engine.resnet_spec()
    p_off, n_param = engine.param_offsets(spec)
    buckets, ready = engine.plan_grad_buckets(spec, 25 * 2 ** 20)
    g = torch.Generator().manual_seed(100 + rank)
    G = torch.randn(n_param, generator=g)
    mine = G.clone()
    # the trainer's loop: after backward op `pos`, all-reduce the bucket it completes
    nops = len(engine.backward_groups(spec))
    for pos in range(nops):
        bi = ready.get(pos)
        if bi is not None:
            a, b = buckets[bi]
            dist.all_reduce(G[a:b])
    G *= 1.0 / world                                   # folded into the SGD kernel on the GPU (hyper[3])
    both = [torch.zeros(n_param) for _ in range(world)]
    dist.all_gather(both, mine)
    expect = sum(both) / world
    ok_avg = torch.allclose(G, expect, atol=1e-6)
    idx = data.distributed_indices(50000, world, rank)
    gathered = [None] * world
    dist.all_gather_object(gathered, idx.tolist())
    if rank == 0:
        allidx = np.concatenate([np.array(x) for x in gathered])
        out.put(dict(ok_avg=bool(ok_avg), n=len(allidx), uniq=len(np.unique(allidx)), buckets=buckets,
                     cover=sum(b - a for a, b in buckets), n_param=n_param))
    dist.barrier()
    dist.destroy_process_group()


def test_bucketed_allreduce_and_sharding_world2():
    world = 2
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, out)) for r in range(world)]
    for p in procs:
        p.start()
    res = out.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res["ok_avg"]
    assert res["n"] == 50000 and res["uniq"] == 50000           # disjoint shards covering the dataset
    assert res["cover"] == res["n_param"]                        # buckets tile the whole arena
    b = res["buckets"]
    assert all(b[i][0] == b[i + 1][1] for i in range(len(b) - 1)) and b[0][1] == res["n_param"] and b[-1][0] == 0
    assert len(b) == 2                                           # 44.7 MB of fp32 grads at 25 MiB per bucket


def test_bucket_plan_small_buckets():
    spec = engine.resnet_spec()
    buckets, ready = engine.plan_grad_buckets(spec, 1 << 20)
    _, n_param = engine.param_offsets(spec)
    assert sum(b - a for a, b in buckets) == n_param
    assert sorted(ready.values()) == list(range(len(buckets)))
    assert max(ready) == len(engine.backward_groups(spec)) - 1  # the last bucket closes after the stem
