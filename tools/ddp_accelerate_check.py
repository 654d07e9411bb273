"""accelerate(net) under DistributedDataParallel on N GPUs (torch.distributed.run, one rank per GPU, NCCL):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
        tools/ddp_accelerate_check.py

Every rank builds the reference's ResNet-18 from the same seed, accelerates it and wraps it as main_dist.py:141-144
does; ranks feed DIFFERENT batches. Checked on every rank, for both ways of doing the all-reduce (attach_ddp():
arena buckets reduced inside backward; plain DDP: the Reducer's own buckets):
  * after backward, every .grad equals the mean over ranks of the per-rank local gradients (computed without DDP
    on a twin network and all-gathered);
  * after optimizer steps the replicas hold identical parameters;
  * BatchNorm running statistics stay per replica during training and follow rank 0 after DDP's
    buffer broadcast at the next forward.
Rank 0 prints one JSON line and writes gpurun_out/ddp_accelerate_check.json."""
import copy
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
import torch.nn as nn

import b200cifar
from b200cifar.accelerate import accelerate, attach_ddp
from oracle.stage_reference import reference_dir


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    ref = reference_dir()
    if ref is not None:
        sys.path.insert(0, ref)
        import models.resnet as R
        make = R.ResNet18
    else:
        from oracle import ref_modules
        make = ref_modules.ResNet18
    N = 128
    rs = np.random.RandomState(100 + rank)                   # a different batch on every rank
    x = torch.from_numpy(rs.standard_normal((N, 3, 32, 32)).astype(np.float32)).to(dev)
    t = torch.from_numpy(rs.randint(0, 10, size=(N,)).astype(np.int64)).to(dev)
    crit = nn.CrossEntropyLoss()
    out = {}
    for mode in ("attached", "reducer"):
        torch.manual_seed(0)
        base = make().to(dev)
        twin = accelerate(copy.deepcopy(base))               # local gradients, no DDP
        crit(twin(x), t).backward()
        local_g = torch.cat([p.grad.reshape(-1) for p in twin.parameters()])
        allg = [torch.empty_like(local_g) for _ in range(world)]
        dist.all_gather(allg, local_g)
        mean_g = torch.stack([g.double() for g in allg]).mean(0)
        net = accelerate(copy.deepcopy(base))
        ddp = nn.parallel.DistributedDataParallel(net, device_ids=[local])      # main_dist.py:141-144
        if mode == "attached":
            attach_ddp(ddp)
        opt = torch.optim.SGD(ddp.parameters(), lr=0.02, momentum=0.9, weight_decay=5e-4)
        opt.zero_grad()
        loss = crit(ddp(x), t)
        loss.backward()
        got = torch.cat([p.grad.reshape(-1) for p in ddp.parameters()]).double()
        err = ((got - mean_g).norm() / mean_g.norm()).item()
        differ = (allg[0] - allg[-1]).abs().max().item() > 0
        opt.step()
        for _ in range(2):
            opt.zero_grad()
            crit(ddp(x), t).backward()
            opt.step()
        flat = torch.cat([p.detach().reshape(-1) for p in ddp.parameters()])
        chk = torch.stack([flat.double().sum(), flat.double().abs().sum()])
        allchk = [torch.zeros_like(chk) for _ in range(world)]
        dist.all_gather(allchk, chk)
        in_sync = all(torch.equal(allchk[0], c) for c in allchk)
        rm = dict(ddp.module.named_buffers())["layer1.0.bn1.running_mean"].clone()
        allrm = [torch.zeros_like(rm) for _ in range(world)]
        dist.all_gather(allrm, rm)
        # DDP broadcasts buffers from rank 0 at the START of each forward; the statistics of that forward itself are
        # per replica (different data), so after a training step they differ across ranks
        bn_local = (allrm[0] - allrm[-1]).abs().max().item() > 0
        out[mode] = dict(grad_rel_err_vs_mean=err, ranks_saw_different_gradients=bool(differ),
                         replicas_in_sync=bool(in_sync), bn_stats_per_replica=bool(bn_local), loss=loss.item())
        ok = err < 1e-4 and differ and in_sync
        flag = torch.tensor([1.0 if ok else 0.0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        out[mode]["ok_all_ranks"] = bool(flag.item() == 1.0)
        del ddp, net, twin
    if rank == 0:
        out["world"] = world
        print(json.dumps(out), flush=True)
        os.makedirs("gpurun_out", exist_ok=True)
        json.dump(out, open("gpurun_out/ddp_accelerate_check.json", "w"), indent=1)
    dist.barrier()
    dist.destroy_process_group()
    if not all(out[m]["ok_all_ranks"] for m in ("attached", "reducer")):
        raise SystemExit(1)


if __name__ == "__main__":
    main()
