"""Which part of the end-to-end loop costs throughput: staggering, read-back, upload?"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import xrtailor_b200 as X
from xrtailor_b200 import scene
sc, body, bindings = scene.body_scene()
seq = np.stack([body.pose(f + 1, amp_deg=30.0) for f in range(60)])
NB, per = int(sys.argv[1]), int(sys.argv[2])
W = int(sys.argv[3]) if len(sys.argv) > 3 else 3
G = int(sys.argv[4]) if len(sys.argv) > 4 else 1
solvers = []
for i in range(NB * per):
    s = scene.make_solver(sc, persist_cluster=1, bindings=bindings)
    s.SetObstacleSequence(seq)
    solvers.append(s)
batches = [X.Batch(solvers[k * per:(k + 1) * per], groups=G, workers=W) for k in range(NB)]
V, Vo = len(sc.cloth_pos), len(sc.obs_pos)
xyz = [np.empty((per * V, 3), np.float32) for _ in batches]
nrm = [np.empty((per * V, 3), np.float32) for _ in batches]
tgt = np.stack([np.concatenate([seq[f]] * per) for f in range(60)])
print('groups', G)
steps = 12
for mode in ["sync", "get", "set", "get+set"]:
    for b in batches:
        b.run(1)
    t0 = time.perf_counter()
    wait = 0.0
    for f in range(steps + 1):
        for k, b in enumerate(batches):
            if f > 0:
                tw = time.perf_counter()
                if "get" in mode:
                    b.GetPositions(0, xyz[k], nrm[k])
                else:
                    b.Sync()
                wait += time.perf_counter() - tw
            if f < steps:
                if "set" in mode:
                    b.UpdateObstacleAnimationFrames(tgt[f % 60])
                b.SimulateSwiftAsync()
    dt = time.perf_counter() - t0
    print(f"{NB} batches x {per}, {W} workers, {mode:8s}: {NB * per * steps / dt:7.0f} frames/s, {dt / steps * 1e3:.1f} ms per step, "
          f"host blocked {wait / steps * 1e3:.1f} ms per step", flush=True)
