"""oracle/ — TEST INFRASTRUCTURE ONLY.

CPU restatements of the reference's algorithm for the hot path (aqualovers/pytorch-cifar). Only
`tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` leg may import
anything from here, and only as the checker or the timed CPU baseline — never as the product path.

Where the arithmetic lives: the reference delegates every operation on this path to third-party
wheels it does not pin (torch 2.11.0+cu128 and torchvision 0.26.0+cu128 are what is installed):
nn.Conv2d / nn.BatchNorm2d / F.relu / F.avg_pool2d / nn.Linear / nn.CrossEntropyLoss / optim.SGD
(models/resnet.py:23-51,102-140; main.py:86-89,99-105) and torchvision.transforms (main.py:30-35).
The restatements therefore are:
  * nets.py     — the reference's model compositions re-expressed functionally on torch fp32 CPU ops
                  (floating-point kernels keep a torch fp32 reference), same construction order so that
                  `torch.manual_seed(s)` yields the reference's initial weights bit for bit;
  * augment.py  — numpy/int restatement of RandomCrop(32, 4) + RandomHorizontalFlip + ToTensor + Normalize
                  including torch's CPU MT19937 draw order (bit-exact);
  * sgd.py      — numpy restatement of optim.SGD(momentum, weight_decay) and CosineAnnealingLR.

Parity pinning: the reference ships no tests or golden vectors, so the pins are outputs of
the reference itself, generated in the build container by tests/golden/make_golden.py (which imports
/root/reference/models and torchvision) and committed under tests/golden/. tests/test_oracle.py checks
every oracle function against those fixtures on CPU.
"""
