"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list: python tools/launch_list.py launches.csv"""
import csv, sys, collections
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot, cnt = collections.Counter(), collections.Counter()
for r in rows[1:]:
    v = float(r[vi].replace(",", ""))
    v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[ui], 1e-6)
    tot[r[ki]] += v
    cnt[r[ki]] += 1
total = sum(tot.values())
print(f"# {sum(cnt.values())} launches, {total:.1f} ms total")
print("# kernel | launches | total ms | share")
for k, v in tot.most_common():
    print(f"{k[:70]:70s} {cnt[k]:7d} {v:9.3f} {100 * v / total:5.1f}%")
