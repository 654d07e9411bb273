"""Attribution experiment for conv_halo_kernel: same launch with parts switched off (desc.flags debug bits)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200cifar
from b200cifar import ops
N = 1024
flush = torch.empty(256 * 2 ** 20, dtype=torch.uint8, device="cuda")
def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize(); ts = []
    for _ in range(reps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b) * 1e3)
    ts.sort(); return ts[len(ts) // 2]
for (hw, cin, cout, bn) in [(32, 64, 64, 0), (16, 128, 128, 0), (16, 128, 128, 64), (8, 256, 256, 0), (8, 256, 256, 256)]:
    x = torch.randn(N, hw, hw, cin, device="cuda").to(torch.bfloat16)
    y = torch.empty(N, hw, hw, cout, device="cuda", dtype=torch.bfloat16)
    wf = ops.pack_weight_fprop(torch.randn(cout, cin, 3, 3, device="cuda"))
    ps = torch.zeros(148, cout, device="cuda"); pq = torch.zeros(148, cout, device="cuda")
    res = {}
    for name, fl in [("full", 0), ("no_stats", 4), ("no_store", 2), ("no_store_no_stats", 6), ("no_tmem_read", 16 + 6),
                     ("no_mma", 8), ("no_mma_no_epi", 8 + 16 + 6), ("nostat_kernel", -1)]:
        if fl == -1:
            d = ops.conv_desc(N, hw, hw, cin, cout, 3, 1, block_n=bn)
            res[name] = round(timeit(lambda: ops.conv_fprop(d, x, wf, y)), 1)
        else:
            d = ops.conv_desc(N, hw, hw, cin, cout, 3, 1, block_n=bn, flags=fl)
            res[name] = round(timeit(lambda: ops.conv_fprop(d, x, wf, y, ps, pq)), 1)
    print(hw, cin, cout, "bn", bn, res, flush=True)
