"""numpy restatement of the reference's training input pipeline (test oracle; bit-exact).

main.py:30-35: Compose([RandomCrop(32, padding=4), RandomHorizontalFlip(), ToTensor(), Normalize(mean, std)])
applied per sample by torchvision 0.26 (not vendored by the reference):
  RandomCrop.forward   pads the uint8 image with 0, then get_params draws i = randint(0, 9), j = randint(0, 9)
  RandomHorizontalFlip draws rand(1) < 0.5
  ToTensor             HWC uint8 -> CHW float32 / 255
  Normalize            (x - mean) / std in float32
Draws come from torch's CPU generator: a 32-bit MT19937; randint(0, n) = raw % n, rand() = (raw & (2^24-1)) * 2^-24.
"""
import numpy as np

MEAN = np.array([0.4914, 0.4822, 0.4465], dtype=np.float32)   # main.py:34
STD = np.array([0.2023, 0.1994, 0.2010], dtype=np.float32)    # main.py:34


class MT19937:
    """Reference 32-bit Mersenne Twister (init_genrand seeding), as used by torch's CPUGeneratorImpl."""

    def __init__(self, seed):
        self.mt = [0] * 624
        self.mt[0] = seed & 0xFFFFFFFF
        for i in range(1, 624):
            self.mt[i] = (1812433253 * (self.mt[i - 1] ^ (self.mt[i - 1] >> 30)) + i) & 0xFFFFFFFF
        self.idx = 624

    def _twist(self):
        mt = self.mt
        for i in range(624):
            y = (mt[i] & 0x80000000) | (mt[(i + 1) % 624] & 0x7FFFFFFF)
            v = mt[(i + 397) % 624] ^ (y >> 1)
            if y & 1:
                v ^= 0x9908B0DF
            mt[i] = v
        self.idx = 0

    def next_u32(self):
        if self.idx >= 624:
            self._twist()
        y = self.mt[self.idx]
        self.idx += 1
        y ^= y >> 11
        y ^= (y << 7) & 0x9D2C5680
        y ^= (y << 15) & 0xEFC60000
        y ^= y >> 18
        return y & 0xFFFFFFFF


def draw_params(rng, n):
    """(i, j, flip) for n consecutive samples, in the order torchvision draws them."""
    out = np.zeros((n, 3), dtype=np.int8)
    for k in range(n):
        out[k, 0] = rng.next_u32() % 9
        out[k, 1] = rng.next_u32() % 9
        u = np.float32(rng.next_u32() & 0xFFFFFF) * np.float32(2.0 ** -24)
        out[k, 2] = 1 if u < np.float32(0.5) else 0
    return out


def augment(images_u8, ijf):
    """images_u8: [n,32,32,3] uint8 HWC; ijf: [n,3]. Returns [n,3,32,32] float32 like the reference's batch."""
    n = images_u8.shape[0]
    out = np.empty((n, 3, 32, 32), dtype=np.float32)
    for k in range(n):
        i, j, flip = int(ijf[k, 0]), int(ijf[k, 1]), int(ijf[k, 2])
        padded = np.zeros((40, 40, 3), dtype=np.uint8)
        padded[4:36, 4:36] = images_u8[k]
        crop = padded[i:i + 32, j:j + 32]
        if flip:
            crop = crop[:, ::-1]
        t = crop.astype(np.float32).transpose(2, 0, 1) / np.float32(255.0)
        out[k] = (t - MEAN[:, None, None]) / STD[:, None, None]
    return out


def normalize_only(images_u8):
    """transform_test (main.py:37-40): ToTensor + Normalize."""
    t = images_u8.astype(np.float32).transpose(0, 3, 1, 2) / np.float32(255.0)
    return (t - MEAN[None, :, None, None]) / STD[None, :, None, None]
