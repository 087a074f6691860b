"""Face-face broad phase (row a16) at C1m / C100k sizes: pairs, device time, rate."""
import json
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import xrtailor_b200 as X
from xrtailor_b200 import scene

out = {}
for name, n, sph in [("C1m", 224, (64, 128)), ("C100k", 317, (64, 128))]:
    pl, tl = scene.grid_cloth(n, 1.0, 0.3, 1e-4)
    pr, tr = scene.uv_sphere(0.3, *sph)
    best = {}
    for mode, args in [("cloth_vs_obstacle", (pr, tr)), ("self", (None, None))]:
        ts = []
        for _ in range(5):
            p, ms = X.traverse_pairs(pl, tl, *args, thickness=2e-3, fast_div=1, timing=True)
            ts.append(ms)
        best[mode] = dict(pairs=int(len(p)), ms=float(np.median(ts)), faces=int(len(tl)),
                          mpairs_per_s=float(len(p) / np.median(ts) / 1e3))
    out[name] = best
print(json.dumps(out, indent=1))

# row a25: FEM strain + isometric bending, ms per Quality-mode iteration
fem = {}
for name, n in [("C1m", 224), ("C100k", 317)]:
    pos, tri = scene.grid_cloth(n, 1.0, 0.0, 1e-4)
    bend_idx, _ = scene.bend_constraints(pos, tri)
    cs, ncs = X.greedy_color(tri, len(pos))
    cb, ncb = X.greedy_color(bend_idx, len(pos))
    x = pos * np.float32(1.05)
    w = np.ones(len(pos), np.float32)
    m = np.array([5000.0, 5000.0, 1000.0, 0.3, 0.25, 0.5], np.float32)
    _, ms = X.fem_solve(pos, x, w, tri, bend_idx, m, cs, cb, 50, 1.0 / 60, timing=True)
    fem[name] = dict(faces=int(len(tri)), hinges=int(len(bend_idx)), strain_colors=int(ncs), bend_colors=int(ncb),
                     ms_per_iteration=float(ms), bytes_per_iteration=int(len(tri) * (36 + 96) + len(bend_idx) * (84 + 128)))
print(json.dumps({"fem": fem}, indent=1))
