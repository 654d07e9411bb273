"""Host side of the input pipeline: the random draws and index lists the reference gets from torch's CPU
generator, reproduced bit for bit so that the CUDA augmentation kernel sees the same (sample, crop, flip)
sequence as `DataLoader(... transform_train ...)` in main.py:30-45 / main_dist.py:93-127.

The pixel work itself runs on the GPU (ops.augment_*); this module is integer bookkeeping in numpy.
"""
import numpy as np

_U32 = np.uint32


class TorchCpuMT19937:
    """32-bit Mersenne Twister as seeded by torch.manual_seed / Generator.manual_seed (init_genrand on the low
    32 bits of the seed), with the block twist vectorised in numpy."""

    def __init__(self, seed):
        mt = np.empty(624, dtype=np.uint64)
        x = int(seed) & 0xFFFFFFFF
        mt[0] = x
        for i in range(1, 624):
            x = (1812433253 * (x ^ (x >> 30)) + i) & 0xFFFFFFFF
            mt[i] = x
        self.mt = mt.astype(_U32)
        self.buf = np.empty(0, dtype=_U32)

    @staticmethod
    def _mix(cur, nxt, far):
        y = (cur & _U32(0x80000000)) | (nxt & _U32(0x7FFFFFFF))
        return far ^ (y >> _U32(1)) ^ ((y & _U32(1)) * _U32(0x9908B0DF))

    def _twist(self):
        mt = self.mt
        mt[0:227] = self._mix(mt[0:227], mt[1:228], mt[397:624])
        mt[227:454] = self._mix(mt[227:454], mt[228:455], mt[0:227])
        mt[454:623] = self._mix(mt[454:623], mt[455:624], mt[227:396])
        mt[623:624] = self._mix(mt[623:624], mt[0:1], mt[396:397])
        y = mt.copy()
        y ^= y >> _U32(11)
        y ^= (y << _U32(7)) & _U32(0x9D2C5680)
        y ^= (y << _U32(15)) & _U32(0xEFC60000)
        y ^= y >> _U32(18)
        return y

    def raw(self, n):
        """Next n raw 32-bit outputs."""
        while self.buf.size < n:
            self.buf = np.concatenate([self.buf, self._twist()])
        out, self.buf = self.buf[:n], self.buf[n:]
        return out


def draw_crop_flip(rng, n):
    """(i, j, flip) int8 [n,3] for n consecutive samples: RandomCrop.get_params draws i then j with
    randint(0, 9) = raw % 9, then RandomHorizontalFlip draws rand(1) = (raw & (2^24-1)) * 2^-24 < 0.5."""
    r = rng.raw(3 * n).reshape(n, 3)
    out = np.empty((n, 3), dtype=np.int8)
    out[:, 0] = r[:, 0] % _U32(9)
    out[:, 1] = r[:, 1] % _U32(9)
    out[:, 2] = (r[:, 2] & _U32(0xFFFFFF)) < _U32(1 << 23)
    return out


def randperm(rng, n):
    """torch.randperm(n, generator=g) on CPU for n < 2^32/20: forward Fisher-Yates with raw % (n - i)."""
    raws = rng.raw(n - 1).astype(np.int64) if n > 1 else np.empty(0, dtype=np.int64)
    z = raws % (n - np.arange(n - 1, dtype=np.int64)) + np.arange(n - 1, dtype=np.int64)
    perm = list(range(n))
    for i, zi in enumerate(z.tolist()):
        perm[i], perm[zi] = perm[zi], perm[i]
    return np.asarray(perm, dtype=np.int64)


def distributed_indices(n, world, rank, seed=0, epoch=0):
    """DistributedSampler(dataset, num_replicas=world, rank=rank) with shuffle=True, drop_last=False
    (main_dist.py:110; torch/utils/data/distributed.py:109-134). The reference never calls set_epoch, so
    epoch stays 0 and every epoch replays the same permutation."""
    perm = randperm(TorchCpuMT19937(seed + epoch), n)
    num_samples = -(-n // world)
    total = num_samples * world
    pad = total - n
    if pad > 0:
        reps = -(-pad // n)
        perm = np.concatenate([perm, np.tile(perm, reps)[:pad]])
    return perm[rank:total:world]


def batches(indices, batch_size):
    """DataLoader(batch_size, drop_last=False) batching: the last batch may be ragged (main.py:44-45)."""
    return [indices[i:i + batch_size] for i in range(0, len(indices), batch_size)]


def synthetic_cifar(n=50000, seed=0):
    """Synthetic stand-in for CIFAR-10 (no network): uint8 [n,32,32,3] images and int64 labels,
    SURVEY.md §8d config 1."""
    rs = np.random.RandomState(seed)
    images = rs.randint(0, 256, size=(n, 32, 32, 3)).astype(np.uint8)
    labels = rs.randint(0, 10, size=(n,)).astype(np.int64)
    return images, labels


def learnable_cifar(n=10000, seed=0, noise=48.0):
    """Synthetic but learnable dataset for accuracy-curve parity: 10 fixed class prototypes + noise."""
    rs = np.random.RandomState(seed)
    protos = rs.randint(32, 224, size=(10, 32, 32, 3)).astype(np.float32)
    labels = rs.randint(0, 10, size=(n,)).astype(np.int64)
    images = protos[labels] + rs.standard_normal((n, 32, 32, 3)).astype(np.float32) * noise
    return np.clip(images, 0, 255).astype(np.uint8), labels


def hard_cifar(n=10000, seed=0, amplitude=6.0, noise=64.0):
    """A synthetic task that does NOT saturate: 10 classes whose prototypes are weak, spatially smooth colour fields
    (a 4x4x3 grid of N(0, amplitude^2) values around mid-grey, each cell 8x8 pixels, so random crops keep most of the
    pattern) buried in per-pixel Gaussian noise. The accuracy a network reaches depends on how well it averages the
    noise out, which makes the top-1 curve sensitive to the numerics of training (tests/test_accuracy_curve_gpu.py).
    The prototypes depend on `seed // 1000` only, so train / test splits drawn with seeds 1000*k + i share them."""
    protos = np.random.RandomState(7919 + seed // 1000).standard_normal((10, 4, 4, 3)).astype(np.float32) * amplitude
    protos = np.repeat(np.repeat(protos, 8, axis=1), 8, axis=2) + 128.0
    rs = np.random.RandomState(seed)
    labels = rs.randint(0, 10, size=(n,)).astype(np.int64)
    images = protos[labels] + rs.standard_normal((n, 32, 32, 3)).astype(np.float32) * noise
    return np.clip(np.rint(images), 0, 255).astype(np.uint8), labels
