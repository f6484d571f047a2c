"""Markdown table of the key ncu counters of every launch of one bench step, from the `--page raw --csv` export that
tools/gpu_profile_full.sh writes:   python tools/ncu_step_table.py gpurun_out/step_full_raw.csv > profiles/rNN_ncu_step_full.md
Launches of the same kernel are averaged.  Times under ncu are cold-cache and serialised: shares, not absolutes."""
import csv, re, sys
from collections import OrderedDict
sys.path.insert(0, __file__.rsplit("/", 2)[0])
from tools.ncu_traffic import tag_of

rows = list(csv.reader(l for l in open(sys.argv[1]) if not l.startswith("==")))
hdr, units, body = rows[0], rows[1], rows[2:]
col = {h: i for i, h in enumerate(hdr)}
UNIT = {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
COLS = [("ms", "gpu__time_duration.sum", 1e3), ("DRAM GB", None, 0),
        ("dram %", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", 1),
        ("L2 %", "lts__throughput.avg.pct_of_peak_sustained_elapsed", 1),
        ("tensor %", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", 1),
        ("fma %", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", 1),
        ("alu %", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", 1),
        ("issue %", "smsp__issue_active.avg.pct_of_peak_sustained_active", 1),
        ("smem pipe %", "l1tex__data_pipe_lsu_wavefronts.sum.pct_of_peak_sustained_elapsed", 1),
        ("warps %", "sm__warps_active.avg.pct_of_peak_sustained_active", 1),
        ("regs", "launch__registers_per_thread", 1), ("grid", "launch__grid_size", 1), ("block", "launch__block_size", 1)]
def num(r, name):
    if name not in col: return float("nan")
    try: return float(r[col[name]].replace(",", "")) * UNIT.get(units[col[name]], 1.0)
    except ValueError: return float("nan")
agg = OrderedDict()
for r in body:
    if len(r) < len(hdr): continue
    tag = tag_of(r[col["Kernel Name"]])
    vals = []
    for label, name, scale in COLS:
        if name is None:
            vals.append((num(r, "dram__bytes_read.sum") + num(r, "dram__bytes_write.sum")) / 1e9)
        else:
            vals.append(num(r, name) * scale)
    a = agg.setdefault(tag, [0, [0.0] * len(COLS)])
    a[0] += 1
    a[1] = [x + y for x, y in zip(a[1], vals)]
tot = sum(a[1][0] for a in agg.values())
print("| kernel | launches | share | " + " | ".join(c[0] for c in COLS) + " |")
print("|---|---|---|" + "---|" * len(COLS))
for tag, (n, v) in sorted(agg.items(), key=lambda kv: -kv[1][1][0]):
    m = [x / n for x in v]
    cells = [f"{m[0]:.3f}", f"{m[1]:.3f}"] + [f"{x:.1f}" for x in m[2:10]] + [f"{x:.0f}" for x in m[10:]]
    print(f"| {tag} | {n} | {100 * v[0] / tot:.1f} % | " + " | ".join(cells) + " |")
