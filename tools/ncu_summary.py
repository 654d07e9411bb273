"""Summarise an `ncu --set full` report of one training step (tools/ncu_step.py) per kernel family:
launches, average duration, DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum), L2->SM bytes
(l1tex__m_xbar2l1tex_read_bytes.sum), and every tensor-pipe style counter the report holds, averaged per launch.

    python tools/ncu_summary.py gpurun_out/r2_conv_step.ncu-rep profiles/r2_ncu_conv_kernels.json [batch]
"""
import csv, io, json, re, subprocess, sys
from collections import OrderedDict, defaultdict

rep, out = sys.argv[1], sys.argv[2]
batch = int(sys.argv[3]) if len(sys.argv) > 3 else 1024
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
names, units = rows[hdr], rows[hdr + 1]
col = {n: i for i, n in enumerate(names)}
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "l1tex__m_xbar2l1tex_read_bytes.sum",
        "lts__t_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__cycles_elapsed.max",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__inst_executed.sum"]
tensor_cols = [n for n in names if re.search(r"tensor|umma|utc", n, re.I)]
UNIT_SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9, "usecond": 1e3,
              "msecond": 1e6, "nsecond": 1, "second": 1e9}


def val(r, n):
    if n not in col:
        return None
    s = r[col[n]].replace(",", "")
    try:
        v = float(s)
    except ValueError:
        return None
    return v * UNIT_SCALE.get(units[col[n]], 1)


def family(k):
    m = re.search(r"(conv_halo2_kernel|conv_halo_kernel|tap_gemm2_kernel|tap_gemm_kernel|wgrad_row2_kernel|wgrad_row_kernel|"
                  r"wgrad_halo_kernel|wgrad_halo_reduce_kernel|wgrad_kernel|bn_\w+_kernel|dw_\w+_kernel|\w+_kernel)", k)
    return m.group(1) if m else k[:40]


fam = defaultdict(list)
for r in rows[hdr + 2:]:
    if len(r) == len(names):
        fam[family(r[col["Kernel Name"]])].append(r)
summary = OrderedDict()
for f, rs in sorted(fam.items(), key=lambda kv: -sum(val(r, "gpu__time_duration.sum") or 0 for r in kv[1])):
    n = len(rs)
    def avg(name):
        vs = [val(r, name) for r in rs]
        vs = [v for v in vs if v is not None]
        return sum(vs) / len(vs) if vs else None
    rd, wr = avg("dram__bytes_read.sum"), avg("dram__bytes_write.sum")
    rec = OrderedDict(launches=n, batch=batch, avg_duration_us=(avg("gpu__time_duration.sum") or 0) / 1e3,
                      avg_dram_read_bytes=rd, avg_dram_write_bytes=wr,
                      avg_dram_bytes_per_launch=(rd + wr) if rd is not None and wr is not None else None,
                      avg_l2_to_sm_bytes=avg("l1tex__m_xbar2l1tex_read_bytes.sum"))
    for w in WANT[4:]:
        rec[w] = avg(w)
    rec["tensor_counters"] = OrderedDict((c, avg(c)) for c in tensor_cols)
    rec["instantiations"] = sorted({re.sub(r"\(.*", "", r[col["Kernel Name"]]).replace("void b2c::<unnamed>::", "") for r in rs})
    summary[f] = rec
json.dump(summary, open(out, "w"), indent=1)
for f, rec in summary.items():
    print(f"{f:28s} n={rec['launches']:3d} {rec['avg_duration_us']:8.1f} us  dram {((rec['avg_dram_bytes_per_launch'] or 0) / 1e6):8.1f} MB"
          f"  l2->sm {((rec['avg_l2_to_sm_bytes'] or 0) / 1e6):8.1f} MB")
