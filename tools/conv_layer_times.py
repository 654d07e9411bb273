"""Per-layer conv kernel timing (CUDA events, isolated launches) for every distinct conv of ResNet-18 at a
given batch: fprop / dgrad / wgrad, halo kernel vs per-tap kernel. Measurement tool, not the product path."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200cifar
from b200cifar import ops

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
shapes = [(32, 64, 64, 3, 1), (32, 64, 128, 3, 2), (32, 64, 128, 1, 2), (16, 128, 128, 3, 1), (16, 128, 256, 3, 2),
          (16, 128, 256, 1, 2), (8, 256, 256, 3, 1), (8, 256, 512, 3, 2), (8, 256, 512, 1, 2), (4, 512, 512, 3, 1),
          (32, 64, 64, 1, 1)]
if os.environ.get("ONLY"):                 # e.g. ONLY=1,2 to profile two shapes under ncu
    shapes = [shapes[int(i)] for i in os.environ["ONLY"].split(",")]
flush = torch.empty(256 * 2 ** 20, dtype=torch.uint8, device="cuda")
scratch = torch.zeros(ops.WGRAD_SCRATCH_FLOATS, device="cuda")


def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


out = []
for (hw, cin, cout, k, s) in shapes:
    ho = hw // s
    x = torch.randn(N, hw, hw, cin, device="cuda").to(torch.bfloat16)
    dy = torch.randn(N, ho, ho, cout, device="cuda").to(torch.bfloat16)
    y = torch.empty_like(dy); dx = torch.empty_like(x)
    w = torch.randn(cout, cin, k, k, device="cuda")
    wf, wd = ops.pack_weight_fprop(w), ops.pack_weight_dgrad(w)
    dw = torch.zeros(cout, cin, k, k, device="cuda")
    ps = torch.zeros(148, cout, device="cuda"); pq = torch.zeros(148, cout, device="cuda")
    fl = 2.0 * N * ho * ho * cout * cin * k * k
    row = dict(shape=f"{hw}x{hw} {cin}->{cout} k{k} s{s}", gflop=fl / 1e9)
    for flags in (0, 1):
        for bn in ((0,) if cout < 256 and cin < 256 else (0, 256)):
            d = ops.conv_desc(N, hw, hw, cin, cout, k, s, flags=flags, block_n=bn)
            tag = f"{'tap' if flags else 'auto'}{'_bn256' if bn else ''}"
            try:
                t = timeit(lambda: ops.conv_fprop(d, x, wf, y, ps, pq))
                row[f"fprop_{tag}_us"] = round(t, 1); row[f"fprop_{tag}_tf"] = round(fl / t / 1e6, 0)
                t = timeit(lambda: ops.conv_dgrad(d, dy, wd, dx))
                row[f"dgrad_{tag}_us"] = round(t, 1); row[f"dgrad_{tag}_tf"] = round(fl / t / 1e6, 0)
            except Exception as e:
                row[f"err_{tag}"] = str(e)[:60]
    for bn in ((0,) if cout < 256 and cin < 256 else (0, 128, 256)):      # CTA-pair kernel (flags bit 7)
        dp = ops.conv_desc(N, hw, hw, cin, cout, k, s, flags=128, block_n=bn)
        tag = f"pair{'_bn%d' % bn if bn else ''}"
        try:
            t = timeit(lambda: ops.conv_fprop(dp, x, wf, y, ps, pq))
            row[f"fprop_{tag}_us"] = round(t, 1); row[f"fprop_{tag}_tf"] = round(fl / t / 1e6, 0)
            if s == 1:
                t = timeit(lambda: ops.conv_dgrad(dp, dy, wd, dx))
                row[f"dgrad_{tag}_us"] = round(t, 1); row[f"dgrad_{tag}_tf"] = round(fl / t / 1e6, 0)
        except Exception as e:
            row[f"err_{tag}"] = str(e)[:60]
    d = ops.conv_desc(N, hw, hw, cin, cout, k, s)
    t = timeit(lambda: ops.conv_wgrad(d, x, dy, dw))
    row["wgrad_atomic_us"] = round(t, 1); row["wgrad_atomic_tf"] = round(fl / t / 1e6, 0)
    t = timeit(lambda: ops.conv_wgrad(d, x, dy, dw, scratch=scratch))
    row["wgrad_us"] = round(t, 1); row["wgrad_tf"] = round(fl / t / 1e6, 0)
    dt = ops.conv_desc(N, hw, hw, cin, cout, k, s, flags=32)          # [tap][cin][cout] gradient layout (engine default)
    t = timeit(lambda: ops.conv_wgrad(dt, x, dy, dw, scratch=scratch))
    row["wgrad_tco_us"] = round(t, 1); row["wgrad_tco_tf"] = round(fl / t / 1e6, 0)
    dt2 = ops.conv_desc(N, hw, hw, cin, cout, k, s, flags=32 | 64)    # ... with the per-tap kernel
    t = timeit(lambda: ops.conv_wgrad(dt2, x, dy, dw, scratch=scratch))
    row["wgrad_tco_tap_us"] = round(t, 1)
    dt3 = ops.conv_desc(N, hw, hw, cin, cout, k, s, flags=32 | 128)   # ... with the CTA-pair row kernel
    t = timeit(lambda: ops.conv_wgrad(dt3, x, dy, dw, scratch=scratch))
    row["wgrad_tco_pair_us"] = round(t, 1); row["wgrad_tco_pair_tf"] = round(fl / t / 1e6, 0)
    if os.environ.get("WGRAD_DEBUG"):      # epilogue cost: plain stores (8) / no stores (16) instead of atomics
        for fl_dbg in (8, 16):
            dd = ops.conv_desc(N, hw, hw, cin, cout, k, s, flags=1 | fl_dbg)
            row[f"wgrad_dbg{fl_dbg}_us"] = round(timeit(lambda: ops.conv_wgrad(dd, x, dy, dw)), 1)
    out.append(row)
    print(json.dumps(row), flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/conv_layer_times.json", "w"), indent=1)
