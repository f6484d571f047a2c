#!/usr/bin/env python
"""Summarise `nvcc -Xptxas -v` logs (puflow_b200/_build/*.ptxas.log): kernel, registers, spills, smem."""
import glob, os, re, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rows = []
for f in sorted(glob.glob(os.path.join(root, "puflow_b200", "_build", "*.ptxas.log"))):
    txt = open(f).read()
    for m in re.finditer(r"Compiling entry function '(\S+)'.*?\n.*?\n\s*(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads\n.*?Used (\d+) registers(?:, used \d+ barriers)?(?:, (\d+) bytes smem)?", txt):
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = re.sub(r"\(anonymous namespace\)::", "", name).split("(")[0].replace("void ", "")
        rows.append((os.path.basename(f)[:-10], name, int(m.group(5)), int(m.group(2)), int(m.group(3)), int(m.group(6) or 0)))
print(f"{'file':10s} {'kernel':44s} {'regs':>5s} {'stack':>6s} {'spill':>6s} {'static smem':>11s}")
for r in rows:
    print(f"{r[0]:10s} {r[1]:44s} {r[2]:5d} {r[3]:6d} {r[4]:6d} {r[5]:11d}")
