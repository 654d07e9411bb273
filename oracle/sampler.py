"""Pure-Python restatement of the reference's sample ordering (test oracle).

main_dist.py:110 `DistributedSampler(trainset)`: torch/utils/data/distributed.py:109-134 — g.manual_seed(seed +
epoch) with seed 0 and epoch never advanced (main_dist.py never calls set_epoch), indices = randperm(n, g),
padded by wrapping to a multiple of the world size, then strided by rank.
torch.randperm on CPU (n < 2^32/20) is a forward Fisher-Yates: z = raw % (n - i); swap(r[i], r[i+z]).
"""
from .augment import MT19937


def randperm(seed, n):
    rng = MT19937(seed)
    r = list(range(n))
    for i in range(n - 1):
        z = rng.next_u32() % (n - i)
        r[i], r[i + z] = r[i + z], r[i]
    return r


def distributed_sampler_indices(n, world, rank, seed=0, epoch=0):
    idx = randperm(seed + epoch, n)
    num_samples = -(-n // world)
    total = num_samples * world
    pad = total - len(idx)
    if pad > 0:
        if pad <= len(idx):
            idx += idx[:pad]
        else:
            idx += (idx * (-(-pad // len(idx))))[:pad]
    return idx[rank:total:world]
