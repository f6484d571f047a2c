"""How much the pointnet2_ops rule "skip candidates with |p|^2 <= 1e-3" (upstream furthest_point_sample, which the reference
calls at modules/utils/patch.py:102,156; NOT in the reference's own modules/utils/fps.py:40-73, which this repo and the
oracle follow) would change on the bench workload.  CPU only (oracle C FPS); prints the figures quoted in DESIGN.md.

    python tools/fps_near_origin.py
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import puflow_oracle as po
from tools import synth

rows = []
for b, shape in enumerate(synth.SHAPES + ("uniform",)):
    pc = synth.make_cloud(2048, 2021 + b, shape)[None]
    c = pc.mean(1, keepdims=True)
    q = pc - c
    pcn = q / np.sqrt((q ** 2).sum(-1)).max()
    near = ((pcn ** 2).sum(-1) <= 1e-3)[0]
    seeds = po.fps(pcn.astype(np.float32), 32)[0]
    # merge-FPS candidates live in the same normalised frame; use the input cloud replicated as a stand-in for their density
    sel = po.fps(pcn.astype(np.float32), 512)[0]
    rows.append((shape, int(near.sum()), int(near[seeds].sum()), int(near[sel].sum())))
print("| cloud (2,048 pts, normalised) | points with norm^2 <= 1e-3 | of the 32 seeds | of 512 FPS samples |")
print("|---|---|---|---|")
for r in rows:
    print("| %s | %d | %d | %d |" % r)
