"""C2: one round of M substep kernels side by side (no other kernels) vs M."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import xrtailor_b200 as X
from xrtailor_b200 import scene
cl = int(sys.argv[1]) if len(sys.argv) > 1 else 8
Ms = [int(a) for a in sys.argv[2:]] or [1, 4, 8, 10, 12, 14, 15, 16, 18]
print("persist_info", X.persist_info(cl))
sc, body, bindings = scene.body_scene()
seq = np.stack([body.pose(f + 1, amp_deg=30.0) for f in range(60)])
solvers = []
for i in range(max(Ms)):
    s = scene.make_solver(sc, persist_cluster=cl, bindings=bindings)
    s.SetObstacleSequence(seq)
    solvers.append(s)
X.batch_run(solvers, 5)
for M in Ms:
    ms, b = X.batch_profile_substep(solvers[:M], reps=8)
    print(f"cluster {cl} M={M}: {ms * 1e3:.0f} us per round -> {M / ms / 4 * 1e3:.0f} f/s equivalent")
