"""C2, shared launch: per-phase-kind time inside garment 0's CTA / cluster while the whole batch runs."""
import os, sys
os.environ["XRT_PERSIST_PROF"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import xrtailor_b200 as X
from xrtailor_b200 import scene

NAMES = ["predict", "collide", "apply1", "interp+normals", "refit", "generate", "stretch", "bend", "lra",
         "solve_vertex", "solve_edge", "apply2", "finalize"]
cl = int(sys.argv[1]) if len(sys.argv) > 1 else 1
Ms = [int(a) for a in sys.argv[2:]] or [1, X.persist_info(cl)[1]]
sc, body, bindings = scene.body_scene()
seq = np.stack([body.pose(f + 1, amp_deg=30.0) for f in range(60)])
varied = os.environ.get("VARIED") == "1"  # phase and amplitude of the motion differ per garment, as in bench.py
rng = np.random.default_rng(20261104)
solvers = []
for i in range(max(Ms)):
    s = scene.make_solver(sc, persist_cluster=cl, bindings=bindings)
    if varied:
        ph, amp = float(rng.uniform(0, 2 * np.pi)), float(rng.uniform(20.0, 30.0))
        s.SetObstacleSequence(np.stack([body.pose(f + 1, amp_deg=amp, phase=ph) for f in range(60)]))
    else:
        s.SetObstacleSequence(seq)
    solvers.append(s)
frames = 30
for M in Ms:
    b = X.Batch(solvers[:M], groups=min(4, M), workers=7)
    b.run(3)
    pa0 = np.stack([s.debug("PERSIST_PROF", np.uint64) for s in solvers[:M]])
    p0 = pa0[0]
    ms = b.run(frames)
    pa = (np.stack([s.debug("PERSIST_PROF", np.uint64) for s in solvers[:M]]) - pa0).astype(np.float64) / frames / 1e3
    p = pa[0]
    inside, between = pa[:, 13], pa[:, 14]
    print(f"  per garment, us/frame: inside shared launches min {inside.min():.0f} mean {inside.mean():.0f} max {inside.max():.0f}; "
          f"between launches min {between.min():.0f} mean {between.mean():.0f} max {between.max():.0f}; "
          f"inside + between mean {(inside + between).mean():.0f}")
    per = max(1, M // min(4, M))
    for gi in range(min(4, M)):
        sl = inside[gi * per:(gi + 1) * per]
        print(f"  group {gi}: slowest garment {sl.max():.0f}, fastest {sl.min():.0f} us/frame inside")
    print(f"cluster {cl} M={M}: {ms / frames:.3f} ms per round = {M * frames / ms * 1e3:.0f} f/s; "
          f"garment 0 per-phase-kind us/frame (sum {p[:13].sum():.0f}):")
    print("   " + "  ".join(f"{nm} {v:.0f}" for nm, v in zip(NAMES, p[:13])))
    b.close()
