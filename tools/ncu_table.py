"""Compact markdown table from the text summaries tools/ncu_summary.py wrote: python tools/ncu_table.py gpurun_out/ncu_*.txt"""
import re, sys
KEYS = [("time ms", "gpu__time_duration.sum"), ("dram rd", "dram__bytes_read.sum"), ("dram wr", "dram__bytes_write.sum"),
        ("dram %", "gpu__dram_throughput"), ("lts %", "lts__throughput"), ("sm %", "sm__throughput"),
        ("fma %", "pipe_fma_cycles"), ("alu %", "pipe_alu_cycles"), ("tensor %", "pipe_tensor_cycles"),
        ("issue %", "smsp__issue_active"), ("occ %", "sm__warps_active"), ("regs", "registers_per_thread"),
        ("smem KB", "shared_mem_per_block_dynamic"), ("grid", "launch__grid_size")]
print("| kernel | " + " | ".join(k for k, _ in KEYS) + " | top stalls |")
print("|---|" + "---|" * (len(KEYS) + 1))
for path in sys.argv[1:]:
    cur, seen = None, set()
    rows = []
    for ln in open(path):
        if ln.startswith("----"):
            cur = {"name": ln[5:].strip(), "stall": []}
            rows.append(cur)
        elif cur is not None:
            m = re.match(r"\s+(\S+)\s+(\S+)\s*(.*)", ln)
            if not m:
                continue
            name, val, unit = m.groups()
            if "issue_stalled" in name:
                try:
                    cur["stall"].append((float(val), name.split("issue_stalled_")[1].split("_per_")[0]))
                except ValueError:
                    pass
            for k, pat in KEYS:
                if pat in name and k not in cur:
                    cur[k] = f"{float(val.replace(',', '')):.4g}" + (" " + unit.split("/")[0] if k in ("time ms", "dram rd", "dram wr") else "")
    for r in rows:
        short = re.sub(r"\(.*", "", r["name"]).replace("void <unnamed>::", "").replace("<unnamed>::", "")
        if short in seen:
            continue
        seen.add(short)
        st = ", ".join(f"{n} {v:.2f}" for v, n in sorted(r["stall"], reverse=True)[:3] if n != "selected")
        print(f"| `{short}` | " + " | ".join(r.get(k, "") for k, _ in KEYS) + f" | {st} |")
