"""GPU-side cost of the per-frame upload: Set + Sync, and a frame with sequence vs uploaded targets."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import xrtailor_b200 as X
from xrtailor_b200 import scene
sc, body, bindings = scene.body_scene()
seq = np.stack([body.pose(f + 1, amp_deg=30.0) for f in range(60)])
n = int(sys.argv[1]) if len(sys.argv) > 1 else 37
solvers = []
for i in range(n):
    s = scene.make_solver(sc, persist_cluster=1, bindings=bindings)
    s.SetObstacleSequence(seq)
    solvers.append(s)
b = X.Batch(solvers, groups=1, workers=7)
tgt = np.stack([np.concatenate([seq[f]] * n) for f in range(60)])
b.run(2)
def t(fn, reps=6):
    out = []
    for r in range(reps):
        b.Sync(); t0 = time.perf_counter(); fn(r); b.Sync(); out.append(time.perf_counter() - t0)
    return np.median(out[1:]) * 1e3
print(f"{n} garments: Set+Sync {t(lambda r: b.UpdateObstacleAnimationFrames(tgt[r])):.2f} ms")
print(f"frame from the resident sequence {t(lambda r: b.SimulateSwiftAsync()):.2f} ms")
def up(r):
    b.UpdateObstacleAnimationFrames(tgt[(r + 9) % 60]); b.SimulateSwiftAsync()
print(f"frame with uploaded targets     {t(up):.2f} ms")
print(f"frame from the resident sequence {t(lambda r: b.SimulateSwiftAsync()):.2f} ms")
