"""Per-source-line stall samples of a kernel in an .ncu-rep: python tools/ncu_hot.py file.ncu-rep [topN]"""
import csv, subprocess, sys, io, collections
out = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'source', '--csv', '--print-source=cuda,sass'],
                     capture_output=True, text=True).stdout
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 25
rows = list(csv.reader(io.StringIO(out)))
cur_file = None; hdr = None
agg = collections.OrderedDict()   # (file, line, text) -> [samples, inst]
line_key = None
sass_hot = []
for r in rows:
    if len(r) >= 2 and r[0] == 'File Path': cur_file = r[1].split('/')[-1]; continue
    if len(r) >= 2 and r[0] == 'Function Name': continue
    if r and r[0] == 'Line No': hdr = r; continue
    if hdr is None or len(r) < len(hdr) - 2: continue
    i_s = hdr.index('# Samples'); i_i = hdr.index('Instructions Executed')
    if r[0] != '':
        line_key = (cur_file, r[0], r[1].strip()[:110]); agg.setdefault(line_key, [0, 0])
    else:
        try:
            s = int(r[i_s]); n = int(r[i_i])
        except ValueError: continue
        if line_key: agg[line_key][0] += s; agg[line_key][1] += n
        sass_hot.append((s, n, r[3].strip()[:90], line_key[1] if line_key else '?'))
tot = sum(v[0] for v in agg.values()) or 1
print(f"total samples {tot}")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:topn]:
    print(f"{v[0]:7d} {100*v[0]/tot:5.1f}%  inst {v[1]:10d}  {k[0]}:{k[1]}  {k[2]}")
print("--- hottest SASS")
for s, n, t, ln in sorted(sass_hot, reverse=True)[:topn]:
    print(f"{s:7d} {100*s/tot:5.1f}%  inst {n:10d}  L{ln}  {t}")
