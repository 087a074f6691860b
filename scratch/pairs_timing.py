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
