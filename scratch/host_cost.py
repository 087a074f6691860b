"""Host-side cost of the three batch calls of the end-to-end loop, GPU idle in between."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import xrtailor_b200 as X
from xrtailor_b200 import scene
sc, body, bindings = scene.body_scene()
seq = np.stack([body.pose(f + 1, amp_deg=30.0) for f in range(8)])
n = int(sys.argv[1]) if len(sys.argv) > 1 else 18
solvers = [scene.make_solver(sc, persist_cluster=1, bindings=bindings) for _ in range(n)]
b = X.Batch(solvers, groups=1, workers=3)
V, Vo = len(sc.cloth_pos), len(sc.obs_pos)
xyz, nrm = np.empty((n * V, 3), np.float32), np.empty((n * V, 3), np.float32)
tgt = np.stack([np.concatenate([seq[f]] * n) for f in range(8)])
T = dict(get=[], set=[], enq=[], gpu=[])
for f in range(8):
    t0 = time.perf_counter(); b.GetPositions(0, xyz, nrm); t1 = time.perf_counter()
    b.UpdateObstacleAnimationFrames(tgt[f]); t2 = time.perf_counter()
    b.SimulateSwiftAsync(); t3 = time.perf_counter()
    b.Sync(); t4 = time.perf_counter()
    T["get"].append(t1 - t0); T["set"].append(t2 - t1); T["enq"].append(t3 - t2); T["gpu"].append(t4 - t3)
for k, v in T.items():
    print(f"{k}: {np.median(v[2:]) * 1e3:.2f} ms per batch of {n} ({np.median(v[2:]) / n * 1e6:.0f} us per garment)")
