import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200cifar
from b200cifar import ops
N = 1024
for (hw, cin, cout, bn) in [(32, 64, 64, 0), (16, 128, 128, 0)]:
    x = torch.randn(N, hw, hw, cin, device="cuda").to(torch.bfloat16)
    y = torch.empty(N, hw, hw, cout, device="cuda", dtype=torch.bfloat16)
    wf = ops.pack_weight_fprop(torch.randn(cout, cin, 3, 3, device="cuda"))
    for fl in (32, 32 + 8 + 16 + 6):
        print("shape", hw, cin, cout, "flags", fl, flush=True)
        d = ops.conv_desc(N, hw, hw, cin, cout, 3, 1, block_n=bn, flags=fl)
        for _ in range(2):
            ops.conv_fprop(d, x, wf, y)
            torch.cuda.synchronize()
