"""Candidate-pair statistics of the obstacle queries on the bench scene."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import xrtailor_b200 as X
from xrtailor_b200 import scene
n = int(sys.argv[1]) if len(sys.argv) > 1 else 224
sc = scene.drop_scene(n)
s = scene.make_solver(sc)
for f in range(int(sys.argv[2]) if len(sys.argv) > 2 else 15):
    s.UpdateObstacleAnimationFrame(scene.sphere_motion(sc.obs_pos, f + 1)); s.SimulateSwift()
s.PredictPositions()
for kind, nm in ((0, "VF point"), (1, "RT ray"), (2, "EE edge")):
    off, items = s.candidates(kind)
    c = np.diff(off)
    a = c[c > 0]
    print(f"{nm}: items {len(c)}, with hits {len(a)}, pairs {c.sum()}, mean(active) {a.mean() if len(a) else 0:.1f}, "
          f"p50 {np.percentile(a,50) if len(a) else 0:.0f}, p99 {np.percentile(a,99) if len(a) else 0:.0f}, max {c.max()}")
vf, rt, ee = s.contacts()
print("live: vf", (vf['index']>=0).sum(), "rt", (rt['index']>=0).sum(), "ee", (ee[:,0]<1e6).sum())
