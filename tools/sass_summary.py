"""Per-kernel SASS mnemonic counts of the built library (run here, no GPU):
    python tools/sass_summary.py > profiles/r02_sass_summary.txt
UTCHMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UBLKCP = cp.async.bulk (TMA 1-D), UTMALDG = TMA tensor load,
LDGSTS = cp.async, SYNCS = mbarrier, FHFMA = mixed fp16 x fp16 + fp32 FMA (the hi/lo split), FFMA2 / FADD2 = packed fp32."""
import os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "puflow_b200", "_lib", "libpuflow_b200.so")
WANT = ["UTCHMMA", "LDTM", "STTM", "UTCBAR", "UBLKCP", "UTMALDG", "LDGSTS", "SYNCS", "FHFMA", "FFMA2", "FADD2", "FFMA", "FMNMX",
        "REDUX", "HMMA"]
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
cur, counts = None, {}
for ln in out.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = cur.replace("(anonymous namespace)::", "").replace("void ", "")
        cur = re.sub(r"\(.*", "", cur)
        counts[cur] = {w: 0 for w in WANT}
        continue
    if cur is None:
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", ln)
    if m:
        op = m.group(1)
        for w in WANT:
            if op == w or (w in ("UTCHMMA", "SYNCS", "UTCBAR", "REDUX", "HMMA") and op.startswith(w)):
                counts[cur][w] += 1
print(f"{'kernel':58s} " + " ".join(f"{w:>7s}" for w in WANT))
for k in sorted(counts):
    c = counts[k]
    print(f"{k[:58]:58s} " + " ".join(f"{c[w]:7d}" for w in WANT))
print("\nlibrary: puflow_b200/_lib/libpuflow_b200.so (sm_100a); HMMA (legacy mma.sync) must be 0 everywhere")
